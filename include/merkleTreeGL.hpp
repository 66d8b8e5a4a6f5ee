// Drop-in replacement for the reference's src/starkpil/merkleTree/merkleTreeGL.{hpp,cpp} (class MerkleTreeGL,
// hpp:9-79, cpp:5-44).  Same public data members (FRI reads them directly, src/starkpil/fri/friProve.cpp:225,241),
// constructors and methods, so src/starkpil/starks.hpp:186-190 (`new MerkleTreeGL(NExtended, width, source)`,
// `new MerkleTreeGL(pConstTreeAddress)`) and the commit stages compile unchanged.  merkelize() -- the
// Poseidon-Goldilocks leaf hashes and tree levels, PoseidonGoldilocks::merkletree_avx
// (src/goldilocks/src/poseidon_goldilocks.cpp:266-301) -- runs on the B200 through glprover.h.
// Header-only: link with -lglprover.  No CPU hashing path exists here.
#ifndef MERKLETREEGL
#define MERKLETREEGL

#if defined(__has_include)
#if __has_include("goldilocks_base_field.hpp")
#include "goldilocks_base_field.hpp"
#else
#include "glp_goldilocks_element.hpp"
#endif
#else
#include "goldilocks_base_field.hpp"
#endif
#include "glp_cxx.hpp"
#include <cstdlib>
#include <cstring>

#ifndef HASH_SIZE
#define HASH_SIZE 4 // poseidon_goldilocks.hpp:10
#endif
#define MERKLEHASHGL_ARITY 2

class MerkleTreeGL
{
private:
    static uint64_t *raw(Goldilocks::Element *p) { return reinterpret_cast<uint64_t *>(p); }

public:
    uint64_t height = 0;
    uint64_t width = 0;
    Goldilocks::Element *source = NULL;
    Goldilocks::Element *nodes = NULL;
    bool isSourceAllocated = false;
    bool isNodesAllocated = false;

    MerkleTreeGL() {}

    // view of a serialised tree [width, height, source..., nodes...] (the *.consttree layout, hpp:24-32)
    MerkleTreeGL(Goldilocks::Element *tree)
    {
        width = Goldilocks::toU64(tree[0]);
        height = Goldilocks::toU64(tree[1]);
        source = tree + 2;
        nodes = source + height * width;
    }

    // source aliased (or zero-allocated when NULL); nodes owned (hpp:33-43)
    MerkleTreeGL(uint64_t _height, uint64_t _width, Goldilocks::Element *_source) : height(_height), width(_width), source(_source)
    {
        if (source == NULL)
        {
            const uint64_t words = height * width;
            source = static_cast<Goldilocks::Element *>(calloc(words > 0 ? words : 1, sizeof(Goldilocks::Element)));
            isSourceAllocated = true;
        }
        const uint64_t n = getTreeNumElements();
        nodes = static_cast<Goldilocks::Element *>(calloc(n > 0 ? n : 1, sizeof(Goldilocks::Element)));
        isNodesAllocated = true;
    }
    MerkleTreeGL(const MerkleTreeGL &) = delete;
    MerkleTreeGL &operator=(const MerkleTreeGL &) = delete;

    ~MerkleTreeGL()
    {
        if (isSourceAllocated)
            free(source);
        if (isNodesAllocated)
            free(nodes);
    }

    void copySource(Goldilocks::Element *_source) { std::memcpy(source, _source, height * width * sizeof(Goldilocks::Element)); }

    // nodes <- all leaf hashes, then every level, root last (cpp:37-44 -> poseidon_goldilocks.cpp:266-301)
    void merkelize()
    {
        if (height == 0)
            return;
        if (glp::multi_enabled()) // GLP_DEVICES: leaf rows split over the devices, top levels on the first (glp_multi_merkletree)
            glp::check(glp_multi_merkletree(glp::thread_multi(), raw(nodes), raw(source), width, height, 1), "MerkleTreeGL::merkelize");
        else
            glp::check(glp_merkletree(glp::thread_ctx(), raw(nodes), raw(source), width, height, 1), "MerkleTreeGL::merkelize");
    }

    uint64_t getTreeNumElements() { return glp_tree_num_elements(height); } // (2*height - 1) * HASH_SIZE (hpp:59-62)

    void getRoot(Goldilocks::Element *root) { std::memcpy(root, nodes + (getTreeNumElements() - HASH_SIZE), HASH_SIZE * sizeof(Goldilocks::Element)); }

    // proof = row idx of source, then the sibling at every level from the leaves up (cpp:12-35)
    void getGroupProof(Goldilocks::Element *proof, uint64_t idx)
    {
        glp::check(glp_merkle_group_proof(glp::thread_ctx(), raw(proof), raw(source), raw(nodes), width, height, idx), "MerkleTreeGL::getGroupProof");
    }

    uint64_t MerkleProofSize()
    {
        uint64_t d = 0; // ceil(log2(height)) (hpp:71-78)
        while (((uint64_t)1 << d) < height)
            d++;
        return d;
    }
};

#endif
