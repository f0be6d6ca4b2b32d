#!/bin/sh
# artisoptions.h for MODE=classic_small: classic mode (see classic.sh) on a 16^3 propagation grid with 4000 packets.
# Small enough that snapshots of it are committed as golden fixtures under tests/golden/.
ARTIS_MPKTS=${ARTIS_MPKTS:-4000} ARTIS_NCOORDGRID=${ARTIS_NCOORDGRID:-16} REF=${REF} sh "$(dirname "$0")/classic.sh"
