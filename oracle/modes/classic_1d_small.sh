#!/bin/sh
# artisoptions.h for MODE=classic_1d_small: classic mode (see classic.sh) with the spherical 1D propagation grid
# (GRID_TYPE GRID_SPHERICAL1D, the commented alternative at artisoptions_classic.h:15: one propagation cell per model
# shell) and 4000 packets. Small enough that snapshots of it are committed as golden fixtures under tests/golden/.
ARTIS_MPKTS=${ARTIS_MPKTS:-4000} REF=${REF} sh "$(dirname "$0")/classic.sh" | sed -e "s/#define GRID_TYPE GRID_UNIFORM/#define GRID_TYPE GRID_SPHERICAL1D/"
