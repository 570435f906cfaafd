// Forwarding header: keeps the reference include path (operators/finetune_ops/core/device.h) valid for unchanged callers.
#pragma once
#include "mft/device.h"
