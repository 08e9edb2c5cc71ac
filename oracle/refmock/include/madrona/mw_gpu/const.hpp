/* shim for M/src/render/ecs_system.cpp:11: the one constant the path reads,
 * GPUImplConsts::get().raycastOutputResolution = 32 (src/entry/mgr.cpp:132). */
#pragma once
#include <cstdint>
#include "host_print.hpp"
namespace madrona { namespace mwGPU {
struct GPUImplConsts {
    uint32_t raycastOutputResolution;
    static GPUImplConsts &get() { static GPUImplConsts c{32u}; return c; }
};
} }
