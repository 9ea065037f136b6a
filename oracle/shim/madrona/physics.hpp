// TEST INFRASTRUCTURE (oracle shim): forwards <madrona/physics.hpp> to the mini-madrona core.
#pragma once
#include "shim_core.hpp"
