// Forwarding header: keeps the reference include path (operators/finetune_ops/graph/gpt2_model.h) valid for unchanged callers.
#pragma once
#include "mft/gpt2_model.h"
