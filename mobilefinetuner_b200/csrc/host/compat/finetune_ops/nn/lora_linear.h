// Forwarding header: keeps the reference include path (operators/finetune_ops/nn/lora_linear.h) valid for unchanged callers.
#pragma once
#include "mft/lora_linear.h"
