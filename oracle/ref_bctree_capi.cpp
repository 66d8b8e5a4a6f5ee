// TEST INFRASTRUCTURE ONLY.  C wrapper around the reference's constant-tree builder, compiled from
// tools/starkpil/bctree/build_const_tree.cpp where it lies (oracle/Makefile target `ref`, oracle/_ref/libbctree.so) with
// two header stand-ins (oracle/shim/nlohmann/json.hpp, oracle/shim/merkleTreeBN128.hpp).  Nothing is restated here:
//   ref_bctree_interpolate -> interpolate()     (build_const_tree.cpp:398-531), bctree's own blocked LDE
//   ref_bctree_build       -> buildConstTree()  (:533-600): constant polynomials file + starkStruct json -> *.consttree image + verkey json
#include <cstdint>
#include <string>
#include "goldilocks_base_field.hpp"
void interpolate(Goldilocks::Element *buffSrc, uint64_t nPols, uint64_t nBits, Goldilocks::Element *buffDst, uint64_t nBitsExt);
void buildConstTree(const std::string constFile, const std::string starkStructFile, const std::string constTreeFile, const std::string verKeyFile);
extern "C" {
int ref_bctree_interpolate(uint64_t *dst, uint64_t *src, uint64_t npols, uint64_t nbits, uint64_t nbits_ext)
{
    interpolate((Goldilocks::Element *)src, npols, nbits, (Goldilocks::Element *)dst, nbits_ext);
    return 0;
}
int ref_bctree_build(const char *const_file, const char *stark_struct_file, const char *const_tree_file, const char *ver_key_file)
{
    buildConstTree(const_file, stark_struct_file, const_tree_file, ver_key_file);
    return 0;
}
}
