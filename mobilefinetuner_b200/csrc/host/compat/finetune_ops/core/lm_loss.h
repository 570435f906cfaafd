// Forwarding header: keeps the reference include path (operators/finetune_ops/core/lm_loss.h) valid for unchanged callers.
#pragma once
#include "mft/lm_loss.h"
