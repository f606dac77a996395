// compatibility forwarder (reference includes this path: src/CG.hpp:15-16, src/VectorOperations.hpp:17)
#include "../../sycl/sycl.hpp"
