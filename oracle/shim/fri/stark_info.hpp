// TEST INFRASTRUCTURE ONLY (oracle/): the part of src/starkpil/stark_info.hpp that FRIProve::prove reads -- the FRI step sizes
// and the query count (stark_info.hpp:21-35).  The real header parses zkevm.starkinfo.json with nlohmann-json and pulls in the
// prover's config; friProve.cpp only touches starkInfo.starkStruct.{steps[i].nBits, nQueries}.
#pragma once
#include <nlohmann/json.hpp>
#include <string>
#include <vector>
#include "goldilocks_base_field.hpp"
#include "polinomial.hpp"
#include "merklehash_goldilocks.hpp"
using json = nlohmann::json;
using namespace std;
class StepStruct
{
public:
    uint64_t nBits;
};
class StarkStruct
{
public:
    uint64_t nBits;
    uint64_t nBitsExt;
    uint64_t nQueries;
    string verificationHashType;
    vector<StepStruct> steps;
};
class StarkInfo
{
public:
    StarkStruct starkStruct;
};
