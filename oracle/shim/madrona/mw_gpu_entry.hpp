// TEST INFRASTRUCTURE (oracle shim): forwards <madrona/mw_gpu_entry.hpp> to the mini-madrona core.
#pragma once
#include "shim_core.hpp"
