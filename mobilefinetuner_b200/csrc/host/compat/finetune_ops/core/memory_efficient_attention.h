// Forwarding header: keeps the reference include path (operators/finetune_ops/core/memory_efficient_attention.h) valid for unchanged callers.
#pragma once
#include "mft/memory_efficient_attention.h"
