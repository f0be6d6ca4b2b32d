#!/bin/sh
# artisoptions.h for MODE=classic_vpkt_small: classic mode (see classic.sh; DIPOLE and POL_ON are on there) with virtual
# packets (VPKT_ON, the commented switch at artisoptions_classic.h:68) on a 16^3 propagation grid with 4000 packets.
# Small enough that snapshots of it are committed as golden fixtures under tests/golden/.
ARTIS_MPKTS=${ARTIS_MPKTS:-4000} ARTIS_NCOORDGRID=${ARTIS_NCOORDGRID:-16} REF=${REF} sh "$(dirname "$0")/classic.sh" | sed -e "s|^// #define VPKT_ON|#define VPKT_ON|"
