/* oracle/shim stand-in (TEST INFRASTRUCTURE): fluid.cc includes <openvdb/tools/PointAdvect.h> but uses no symbol
 * that is not already provided by the dense-array stand-in <openvdb/openvdb.h> of this directory. */
#include <openvdb/openvdb.h>
