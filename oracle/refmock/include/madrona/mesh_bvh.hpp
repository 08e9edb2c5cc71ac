/* shim: the sensor path never touches a mesh BVH (agents are spheres); only the names are needed by
 * M/src/mw/device/include/madrona/bvh.hpp:224 and M/src/render/ecs_system.cpp:41. */
#pragma once
#include <cstdint>
#include <madrona/types.hpp>
namespace madrona { namespace render {
struct MeshBVH;
/* M/include/madrona/mesh_bvh_compressed.hpp:16-35 (the TLAS traversal's node stack) */
struct TraversalStack {
    static constexpr CountT stackSize = 32;
    int32_t s[stackSize];
    CountT size;
    void push(int32_t v) { s[size++] = v; }
    int32_t pop() { return s[--size]; }
};
} using MeshBVH = render::MeshBVH; }
typedef unsigned long long cudaTextureObject_t;
