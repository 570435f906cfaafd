// Forwarding header: keeps the reference include path (operators/finetune_ops/core/dtype.h) valid for unchanged callers.
#pragma once
#include "mft/dtype.h"
