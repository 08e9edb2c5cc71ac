/* shim for M/include/madrona/render/ecs.hpp: the declarations src/sim/sim.cpp (HEAD) and the vendored
 * M/src/render/ecs_system.cpp + M/src/mw/device/bvh_raycast.cpp are written against. Field lists follow the
 * vendored header (:13-88, :119-157); added on top, because HEAD's sim.cpp names them and the vendored
 * header predates it: SemanticOutputBuffer / DepthOutputBuffer / Prev*OutputBuffer
 * (src/sim/sim.cpp:98-112,748-762) and InstanceData::speciesIDX (:787-788). */
#pragma once
#include <madrona/math.hpp>
#include <madrona/taskgraph_builder.hpp>
namespace madrona { namespace render {
struct RenderCamera { Entity cameraEntity; float fovScale; float zNear; math::Vector3 cameraOffset; };
struct Renderable { Entity renderEntity; };
using MortonCode = uint32_t;
struct alignas(16) PerspectiveCameraData {
    math::Vector2 position; float viewDirPolar; uint8_t numForwardRays; uint8_t numBackwardRays;
    int32_t worldIDX; uint32_t rowIDX; float zOffset;
};
struct alignas(16) InstanceData {
    math::Vector2 position; math::Vector2 scale; float viewDirPolar; Entity owner;
    int32_t objectIDX; int32_t worldIDX; float zOffset;
    int32_t speciesIDX;                         /* HEAD only */
};
struct RenderOutputBufferImpl { char buffer[1]; };
struct RenderOutputBuffer : RenderOutputBufferImpl {};
struct FinderOutput { Entity hitEntity; float depth; };
struct FinderOutputBuffer : RenderOutputBufferImpl {};
struct SensorRow { char buffer[32]; };          /* raycastOutputResolution bytes (src/entry/mgr.cpp:132) */
struct SemanticOutputBuffer : SensorRow {};
struct DepthOutputBuffer : SensorRow {};
struct PrevSemanticOutputBuffer : SensorRow {};
struct PrevDepthOutputBuffer : SensorRow {};
struct RenderOutputRef { Entity outputEntity; Entity finderOutputEntity; };
struct alignas(16) TLBVHNode { math::AABB aabb; };
struct RenderableArchetype : public Archetype<InstanceData, MortonCode, TLBVHNode> {};
struct RenderCameraArchetype : public Archetype<PerspectiveCameraData, RenderOutputRef> {};
struct RaycastOutputArchetype : public Archetype<RenderOutputBuffer, FinderOutputBuffer, SemanticOutputBuffer,
                                                 DepthOutputBuffer, PrevSemanticOutputBuffer, PrevDepthOutputBuffer> {};
struct RenderECSBridge;
namespace RenderingSystem {
void registerTypes(ECSRegistry &registry, const RenderECSBridge *bridge);
TaskGraphNodeID setupTasks(TaskGraphBuilder &builder, Span<const TaskGraphNodeID> deps);
void init(Context &ctx, const RenderECSBridge *bridge);
uint32_t *getVoxelPtr(Context &ctx);
void makeEntityRenderable(Context &ctx, Entity e);
void attachEntityToView(Context &ctx, Entity e, float vfov_degrees, float z_near, const math::Vector3 &camera_offset);
void cleanupViewingEntity(Context &ctx, Entity e);
void cleanupRenderableEntity(Context &ctx, Entity e);
/* M/include/madrona/render/ecs.inl:5-26 (GPU branch): the output record of a camera is the row of its camera
 * entity in the raycast output table */
template <typename OutputT> OutputT *getRenderOutput(Context &ctx, const RenderCamera &camera, uint32_t output_size)
{
    const Loc cam = ctx.loc(camera.cameraEntity);
    uint8_t *p = (uint8_t *)mwGPU::getStateManager()->getArchetypeComponent<RaycastOutputArchetype, OutputT>();
    return (OutputT *)(p + (size_t)output_size * cam.row);
}
}
} }
