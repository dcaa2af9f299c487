// orc_system.cpp — TEST INFRASTRUCTURE: the same FullSystem control flow (include/ldso_system.hpp, restating
// src/frontend/FullSystem.cc:78-2109 and the host half of EnergyFunctional.cc) bound to the CPU oracle's orc_* entry
// points, exported as orc_system_*.  Lets tests / the CPU baseline run the whole per-frame pipeline on the host and
// compare it call for call with ldso_gpu_system_*.  Never linked into or loaded by the product library.
#include <new>

#include "orc_common.h"
#include "../include/ldso_system.hpp"

LDSO_DECLARE_API(orc_, orc_ctx)
LDSO_DEFINE_API(OrcApi, orc_, orc_ctx)
LDSO_DEFINE_SYSTEM_EXPORTS(orc_, OrcApi)
