#!/bin/sh
# artisoptions.h for MODE=kilonova: the reference's artisoptions_kilonova_lte.h with the overrides of its own test setup
# (tests/setup_kilonova.sh:21-27: TABLESIZE 20, MINTEMP 1000, MAXTEMP 20000, partial emission/absorption spectra);
# packet count and grid size can be overridden.
MPKTS=${ARTIS_MPKTS:-40000}
NGRID=${ARTIS_NCOORDGRID:-50}
sed -e "s/#define MPKTS.*/#define MPKTS ${MPKTS}/" \
    -e "s/#define CUBOID_NCOORDGRID_X 50/#define CUBOID_NCOORDGRID_X ${NGRID}/" \
    -e "s/#define CUBOID_NCOORDGRID_Y 50/#define CUBOID_NCOORDGRID_Y ${NGRID}/" \
    -e "s/#define CUBOID_NCOORDGRID_Z 50/#define CUBOID_NCOORDGRID_Z ${NGRID}/" \
    -e "s/#define TABLESIZE.*/#define TABLESIZE 20/" \
    -e "s/#define MINTEMP.*/#define MINTEMP 1000./" \
    -e "s/#define MAXTEMP.*/#define MAXTEMP 20000./" \
    -e "s/constexpr bool WRITE_PARTIAL_EMISSIONABSORPTIONSPEC.*/constexpr bool WRITE_PARTIAL_EMISSIONABSORPTIONSPEC = true;/" \
    "${REF}/artisoptions_kilonova_lte.h"
