// ORACLE -- TEST INFRASTRUCTURE ONLY: forwards to the GBWTGraph stand-in (utils.h).
#pragma once
#include "utils.h"
