// TEST INFRASTRUCTURE ONLY (oracle/): src/utils/zklog.hpp is the prover's logger (config, mutex, json); polinomial.hpp and
// friProve.cpp only call zklog.error / zklog.info with a string.
#pragma once
#include <cassert>
#include <iostream>
#include <map>
#include <string>
#include <vector>
using namespace std;                // the real header's includes bring this into every file that includes it (polinomial.hpp relies on it)
class zkLog
{
public:
    void error(const std::string &s) { std::cerr << s << std::endl; }
    void info(const std::string &) {}
    void warning(const std::string &s) { std::cerr << s << std::endl; }
};
static zkLog zklog;
#ifndef zkassert
#define zkassert(a) assert(a)       // src/utils/zkassert.hpp
#endif
