// GSL shim forwarding header (test infrastructure; see gslshim_all.h)
#include "gslshim_all.h"
