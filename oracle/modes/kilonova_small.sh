#!/bin/sh
# artisoptions.h for MODE=kilonova_small: kilonova LTE mode (see kilonova.sh) on a 20^3 propagation grid with 4000 packets,
# small enough that snapshots of it are committed as golden fixtures under tests/golden/.
ARTIS_MPKTS=${ARTIS_MPKTS:-4000} ARTIS_NCOORDGRID=${ARTIS_NCOORDGRID:-20} REF=${REF} sh "$(dirname "$0")/kilonova.sh"
