// Forwarding header: keeps the reference include path (operators/finetune_ops/core/backward_functions.h) valid for unchanged callers.
#pragma once
#include "mft/autograd_engine.h"
