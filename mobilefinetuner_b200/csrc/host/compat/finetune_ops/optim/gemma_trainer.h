// Forwarding header: keeps the reference include path valid for unchanged callers.
#pragma once
#include "mft/gemma_trainer.h"
