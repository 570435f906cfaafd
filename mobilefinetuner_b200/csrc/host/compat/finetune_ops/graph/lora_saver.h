// Forwarding header: keeps the reference include path valid for unchanged callers.
#pragma once
#include "mft/lora_saver.h"
