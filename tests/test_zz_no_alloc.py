"""The allocation-free entry point, glp_merkle_group_proofs_static (include/glprover.h): batched Merkle group proofs whose
index staging lives in a file-scope `static __device__` array instead of the context's cudaMalloc'ed scratch.  It has no
counterpart in the reference; it is checked, through the C ABI, against the existing glp_merkle_group_proofs on the same
device-resident trees -- bit-exact, like every integer result in this suite -- on two calls in a row per shape (the second
shorter than the first, so stale entries of the shared array would show), at the limit, and past it.

The file is named so that pytest collects it LAST: a failure here must not hide the
parity tests.  It needs nothing from the other tests (no oracle, no fixtures beyond the context), so it passes alone in a
fresh process as well as after them."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_group_proofs_static_matches_allocating_entry_point(ctx):
    from cases import check_group_proofs_static
    rng = np.random.default_rng(20261105)
    check_group_proofs_static(ctx, rng, shapes=((1 << 10, 21), (1 << 14, 40), (64, 128), (2, 5), (1, 9)))


def test_group_proofs_static_second_context_same_device(ctx):
    """a second context on the same device reuses the array (its address is looked up once per device, not per context)"""
    from cases import check_group_proofs_static
    from xlayer_prover_b200 import Context
    c2 = Context(0)
    try:
        check_group_proofs_static(c2, np.random.default_rng(7), shapes=((256, 8),))
    finally:
        c2.close()
    check_group_proofs_static(ctx, np.random.default_rng(8), shapes=((256, 8),))
