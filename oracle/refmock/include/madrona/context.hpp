/* shim for src/sim/sim.hpp:8: CustomContext<Engine, Sim> comes from the mock (oracle/refmock/mock_ecs.hpp). */
#pragma once
#include "../../mock_ecs.hpp"
