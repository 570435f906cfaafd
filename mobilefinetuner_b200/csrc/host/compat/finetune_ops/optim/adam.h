// Forwarding header: keeps the reference include path (operators/finetune_ops/optim/adam.h) valid for unchanged callers.
#pragma once
#include "mft/adam.h"
