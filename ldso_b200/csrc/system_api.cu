// system_api.cu — ldso_gpu_system_*: FullSystem::addActiveFrame as one C-ABI call per frame (include/ldso_system.h).
// The control flow lives in include/ldso_system.hpp (host C++, the reference's FullSystem.cc / EnergyFunctional.cc
// restated over the C ABI); this translation unit binds it to the ldso_gpu_* entry points of this library.
#include <new>

#include "../../include/ldso_system.hpp"

LDSO_DEFINE_API(LdsoGpuApi, ldso_gpu_, ldso_gpu_ctx)
LDSO_DEFINE_SYSTEM_EXPORTS(ldso_gpu_, LdsoGpuApi)
