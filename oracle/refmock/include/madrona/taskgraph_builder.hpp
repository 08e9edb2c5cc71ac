/* shim: the reference's src/sim/sim.hpp:9 and M/include/madrona/render/ecs.hpp:4 include this for
 * TaskGraphBuilder / ParallelForNode / SortArchetypeNode; the mock provides them (oracle/refmock/mock_ecs.hpp). */
#pragma once
#include "../../mock_ecs.hpp"
