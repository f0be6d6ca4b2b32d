#!/bin/sh
# artisoptions.h for MODE=classic_vpkt: BASELINE.json config 5 -- classic mode (see classic.sh; DIPOLE and POL_ON are on
# there) with virtual packets (VPKT_ON, the commented switch at artisoptions_classic.h:68) on the full 100^3 grid.
REF=${REF} sh "$(dirname "$0")/classic.sh" | sed -e "s|^// #define VPKT_ON|#define VPKT_ON|"
