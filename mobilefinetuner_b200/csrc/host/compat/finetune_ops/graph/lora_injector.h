// Forwarding header: keeps the reference include path (operators/finetune_ops/graph/lora_injector.h) valid for unchanged callers.
#pragma once
#include "mft/lora_injector.h"
