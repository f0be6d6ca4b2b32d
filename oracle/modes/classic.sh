#!/bin/sh
# artisoptions.h for MODE=classic: the reference's artisoptions_classic.h with the overrides its own test setup
# applies (tests/setup_classicmode.sh:15-17); MPKTS can be overridden with ARTIS_MPKTS.
MPKTS=${ARTIS_MPKTS:-15000}
NGRID=${ARTIS_NCOORDGRID:-100}
sed -e "s/#define MPKTS 100000/#define MPKTS ${MPKTS}/" \
    -e "s/#define CUBOID_NCOORDGRID_X 100/#define CUBOID_NCOORDGRID_X ${NGRID}/" \
    -e "s/#define CUBOID_NCOORDGRID_Y 100/#define CUBOID_NCOORDGRID_Y ${NGRID}/" \
    -e "s/#define CUBOID_NCOORDGRID_Z 100/#define CUBOID_NCOORDGRID_Z ${NGRID}/" \
    -e "s/WRITE_PARTIAL_EMISSIONABSORPTIONSPEC = false/WRITE_PARTIAL_EMISSIONABSORPTIONSPEC = true/" \
    "${REF}/artisoptions_classic.h"
