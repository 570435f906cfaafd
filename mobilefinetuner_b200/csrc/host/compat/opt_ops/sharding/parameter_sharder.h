// Forwarding header: keeps the reference include path (operators/opt_ops/sharding/parameter_sharder.h) valid for unchanged callers.
#pragma once
#include "mft/parameter_sharder.h"
