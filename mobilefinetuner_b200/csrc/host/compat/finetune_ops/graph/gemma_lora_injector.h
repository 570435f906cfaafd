// Forwarding header: keeps the reference include path (operators/finetune_ops/graph/gemma_lora_injector.h) valid for unchanged callers.
#pragma once
#include "mft/gemma_lora_injector.h"
