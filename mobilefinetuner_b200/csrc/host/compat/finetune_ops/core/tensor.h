// Forwarding header: keeps the reference include path (operators/finetune_ops/core/tensor.h) valid for unchanged callers.
#pragma once
#include "mft/tensor.h"
