// TEST INFRASTRUCTURE (oracle shim): forwards <madrona/registry.hpp> to the mini-madrona core.
#pragma once
#include "shim_core.hpp"
