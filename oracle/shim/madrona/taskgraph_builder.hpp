// TEST INFRASTRUCTURE (oracle shim): forwards <madrona/taskgraph_builder.hpp> to the mini-madrona core.
#pragma once
#include "shim_core.hpp"
