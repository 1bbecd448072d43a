// forwarding header: the reference includes this OpenFHE path; everything lives in openfhe.h (the lbcrypto shim)
#pragma once
#include "../../openfhe.h"
