// Forwarding header: keeps the reference include path (operators/finetune_ops/optim/optimizer.h) valid for unchanged callers.
#pragma once
#include "mft/optimizer.h"
