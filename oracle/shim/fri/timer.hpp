// TEST INFRASTRUCTURE ONLY (oracle/): src/utils/timer.hpp (every use in friProve.cpp is commented out)
#pragma once
#define TimerStart(name)
#define TimerStop(name)
#define TimerLog(name)
#define TimerStopAndLog(name)
