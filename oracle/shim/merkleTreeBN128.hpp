// TEST INFRASTRUCTURE ONLY (oracle/): build_const_tree.cpp also knows a BN128 flavour of the constant tree (Poseidon over the
// BN254 scalar field, src/starkpil/merkleTree/merkleTreeBN128.*, which needs the ffiasm-generated fr.hpp).  The oracle only
// ever builds "GL" trees; this stand-in lets the file compile and aborts if that branch is reached.
#pragma once
#include <cstdint>
#include <cstdlib>
#include <string>
#include "goldilocks_base_field.hpp"
struct RawFr {
    struct Element { uint64_t v[4]; };
    std::string toString(const Element &) { abort(); }
};
class MerkleTreeBN128 {
public:
    uint64_t source_width = 0, height = 0, numNodes = 0;
    Goldilocks::Element *source = nullptr;
    RawFr::Element *nodes = nullptr;
    MerkleTreeBN128(uint64_t, uint64_t, Goldilocks::Element *) { abort(); }
    void merkelize() { abort(); }
    void getRoot(RawFr::Element *) { abort(); }
};
