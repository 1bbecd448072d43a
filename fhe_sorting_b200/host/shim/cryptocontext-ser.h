// forwarding header: the serialisation entry points live in openfhe.h (namespace lbcrypto::Serial)
#pragma once
#include "openfhe.h"
