// TEST INFRASTRUCTURE ONLY. Forwarding header: the reference includes this path
// (src/CG.hpp:16, src/VectorOperations.hpp:17); everything lives in the single stand-in header.
#include "../../../sycl/sycl.hpp"
