// TEST INFRASTRUCTURE (oracle shim): forwards <madrona/navmesh.hpp> to the mini-madrona core.
#pragma once
#include "shim_core.hpp"
