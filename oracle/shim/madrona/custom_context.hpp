// TEST INFRASTRUCTURE (oracle shim): forwards <madrona/custom_context.hpp> to the mini-madrona core.
#pragma once
#include "shim_core.hpp"
