#!/bin/sh
# artisoptions.h for MODE=nebular: the reference's artisoptions_nltenebular.h with the overrides of its own test setup
# (tests/setup_nebularonezone.sh:19-31: TABLESIZE 20, MINTEMP 2000, MAXTEMP 10000, FIRST_NLTE_RADFIELD_TIMESTEP 7,
# DETAILED_BF_ESTIMATORS_USEFROMTIMESTEP 7, SF_AUGER_CONTRIBUTION_ON false, partial emission/absorption spectra);
# packet count, grid size and the Spencer-Fano energy grid can be overridden.
MPKTS=${ARTIS_MPKTS:-1000000}
NGRID=${ARTIS_NCOORDGRID:-50}
SFPTS=${ARTIS_SFPTS:-4096}
sed -e "s/#define MPKTS.*/#define MPKTS ${MPKTS}/" \
    -e "s/#define CUBOID_NCOORDGRID_X 50/#define CUBOID_NCOORDGRID_X ${NGRID}/" \
    -e "s/#define CUBOID_NCOORDGRID_Y 50/#define CUBOID_NCOORDGRID_Y ${NGRID}/" \
    -e "s/#define CUBOID_NCOORDGRID_Z 50/#define CUBOID_NCOORDGRID_Z ${NGRID}/" \
    -e "s/#define TABLESIZE.*/#define TABLESIZE 20/" \
    -e "s/#define MINTEMP.*/#define MINTEMP 2000./" \
    -e "s/#define MAXTEMP.*/#define MAXTEMP 10000./" \
    -e "s/#define SFPTS.*/#define SFPTS ${SFPTS}/" \
    -e "s/constexpr int FIRST_NLTE_RADFIELD_TIMESTEP.*/constexpr int FIRST_NLTE_RADFIELD_TIMESTEP = 7;/" \
    -e "s/#define DETAILED_BF_ESTIMATORS_USEFROMTIMESTEP.*/#define DETAILED_BF_ESTIMATORS_USEFROMTIMESTEP 7/" \
    -e "s/#define SF_AUGER_CONTRIBUTION_ON.*/#define SF_AUGER_CONTRIBUTION_ON false/" \
    -e "s/constexpr bool WRITE_PARTIAL_EMISSIONABSORPTIONSPEC.*/constexpr bool WRITE_PARTIAL_EMISSIONABSORPTIONSPEC = true;/" \
    "${REF}/artisoptions_nltenebular.h"
