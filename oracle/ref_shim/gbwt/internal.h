// ORACLE -- TEST INFRASTRUCTURE ONLY: forwards to the GBWT stand-in (utils.h).
#pragma once
#include "utils.h"
