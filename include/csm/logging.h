/* include/csm/logging.h -- forwarding header: the reference includes <csm/logging.h>
 * (src/zhcsm/ZHCanonical_Matcher.cpp:10-12); with -I<repo>/include it gets the CSM-shaped entry of libzhsm.so. */
#ifndef ZHSM_CSM_FWD_LOGGING_H
#define ZHSM_CSM_FWD_LOGGING_H
#include "../csm_compat.h"
#endif
