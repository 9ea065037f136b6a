// TEST INFRASTRUCTURE (oracle shim): forwards <madrona/taskgraph.hpp> to the mini-madrona core.
#pragma once
#include "shim_core.hpp"
