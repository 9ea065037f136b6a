// TEST INFRASTRUCTURE (oracle shim): forwards <madrona/rand.hpp> to the mini-madrona core.
#pragma once
#include "shim_core.hpp"
