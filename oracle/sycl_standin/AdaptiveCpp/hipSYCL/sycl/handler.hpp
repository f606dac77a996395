// TEST INFRASTRUCTURE ONLY. Forwarding header: the reference includes this path
// (src/CG.hpp:15); everything lives in the single stand-in header.
#include "../../sycl/sycl.hpp"
