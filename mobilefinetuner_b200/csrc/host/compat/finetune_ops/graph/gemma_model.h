// Forwarding header: keeps the reference include path (operators/finetune_ops/graph/gemma_model.h) valid for unchanged callers.
#pragma once
#include "mft/gemma_model.h"
