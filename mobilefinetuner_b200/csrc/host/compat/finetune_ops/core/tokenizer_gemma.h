// Forwarding header: keeps the reference include path valid for unchanged callers.
#pragma once
#include "mft/tokenizer_gemma.h"
