// forwarding header: the reference's generated module loader includes "include/<Impl>.hpp"
#include "../GpuStep.hpp"
