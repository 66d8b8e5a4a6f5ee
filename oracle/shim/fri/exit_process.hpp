// TEST INFRASTRUCTURE ONLY (oracle/): src/utils/exit_process.hpp
#pragma once
#include <cstdlib>
static inline void exitProcess(void) { abort(); }
