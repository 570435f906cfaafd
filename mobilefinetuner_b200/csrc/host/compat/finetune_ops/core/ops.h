// Forwarding header: keeps the reference include path (operators/finetune_ops/core/ops.h) valid for unchanged callers.
#pragma once
#include "mft/ops.h"
