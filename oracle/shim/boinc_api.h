/* TEST INFRASTRUCTURE ONLY (oracle/): erp_utilities.cpp:31 includes <boinc_api.h> for one
 * function (resolveFilename, erp_utilities.cpp:211-214) that is not on the hot path. */
#ifndef BRP_ORACLE_SHIM_BOINC_API_H
#define BRP_ORACLE_SHIM_BOINC_API_H
#include <string.h>
static inline int boinc_resolve_filename(const char *logical, char *physical, int len)
{
    strncpy(physical, logical, (size_t)len);
    if (len > 0) physical[len - 1] = 0;
    return 0;
}
#endif
