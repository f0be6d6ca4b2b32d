#!/bin/sh
# artisoptions.h for MODE=nebular_small: nebular NLTE mode (see nebular.sh) on a 16^3 propagation grid with 20000 packets
# and a 512-point Spencer-Fano energy grid, small enough that snapshots of it are committed as golden fixtures.
ARTIS_MPKTS=${ARTIS_MPKTS:-20000} ARTIS_NCOORDGRID=${ARTIS_NCOORDGRID:-16} ARTIS_SFPTS=${ARTIS_SFPTS:-512} REF=${REF} sh "$(dirname "$0")/nebular.sh"
