// ORACLE -- TEST INFRASTRUCTURE ONLY: forwards to the SDSL stand-in (see ../sdsl_shim.hpp).
#pragma once
#include "../sdsl_shim.hpp"
