/* shim for src/sim/sim.cpp:4: the world-init kernel generator is a no-op on the host
 * (same one-line definition as M/include/madrona/mw_gpu_entry.hpp:3). */
#pragma once
#define MADRONA_BUILD_MWGPU_ENTRY(ContextT, WorldT, ConfigT, InitT)
