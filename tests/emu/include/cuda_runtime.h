// TEST INFRASTRUCTURE ONLY: stand-in for the CUDA runtime header in the emulator build (see poa_emu.hpp).
#pragma once
#include "poa_emu.hpp"
