/* shim for M/include/madrona/components.hpp:8-29. Position / Rotation / Scale are the same thin wrappers over the
 * reference's own math types; ObjectID has the TWO fields src/sim/sim.cpp:213-216 (HEAD) initialises -- the
 * vendored header (one field) is older than the sim. */
#pragma once
#include <madrona/math.hpp>
#include "../../mock_ecs.hpp"
namespace madrona { namespace base {
struct Position : math::Vector3 { Position(math::Vector3 v) : Vector3(v) {} };
struct Rotation : math::Quat { Rotation(math::Quat q) : Quat(q) {} };
struct Scale : math::Diag3x3 { Scale(math::Diag3x3 d) : Diag3x3(d) {} };
struct ObjectID { int32_t idx; int32_t speciesIdx; };
inline void registerTypes(ECSRegistry &) {}
} }
