// TEST INFRASTRUCTURE ONLY.  The translation unit that compiles the reference's tools/starkpil/bctree/build_const_tree.cpp, unmodified,
// into oracle/_ref/libbctree.so.
//
// One thing about the ENVIRONMENT is reproduced: bctree's interpolate() takes the zero padding of the low-degree extension from the
// upper, never written half of buffers it gets from malloc (build_const_tree.cpp:401 tmpBuff, :492-494 bitReverse over all nExt rows;
// buildConstTree's constPolsArrayE, :553).  That holds in the tool's own process -- a fresh multi-gigabyte malloc is zero pages from
// mmap -- and not in a long-lived test process whose heap hands back recycled memory (interpolate then differs from extendPol at
// every size below the mmap threshold and equals it above: tests/test_oracle.py).  So, for this file only, malloc is calloc.
// Every header the file includes is included first, so the macro touches nothing but the file's own calls.
#include <nlohmann/json.hpp>
#include <algorithm>
#include <cstdint>
#include <cstdlib>
#include <filesystem>
#include <fstream>
#include <iostream>
#include <string>
#include <fcntl.h>
#include <immintrin.h>
#include <math.h>
#include <omp.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <sys/time.h>
#include <unistd.h>
#include "goldilocks_base_field.hpp"
#include "merklehash_goldilocks.hpp"
#include "poseidon_goldilocks.hpp"
#include "merkleTreeBN128.hpp"
#include "build_const_tree.hpp"
#define malloc(x) calloc(1, (x))
#include "build_const_tree.cpp"
