"""Own host index builder (SA-IS) vs the reference's `bwa index` arrays: byte-identical."""
import numpy as np
import pytest

import common as cm
from bwa_mem_gpu_b200 import lib as L

needs_ref = pytest.mark.skipif(not cm.have_ref(), reason="oracle/_ref not built")


@needs_ref
@pytest.mark.parametrize("gname", ["tiny", "small", "rep"])
def test_host_index_matches_reference(gname):
    idx = cm.ref_index(gname)
    h = L.HostIndex(cm.genome(gname), cm.backend_lib("emu"))
    assert (h.primary, h.L2, h.seq_len, h.bwt_size, h.n_sa, h.sa_intv) == \
           (idx.primary, idx.L2, idx.seq_len, idx.bwt_size, idx.n_sa, idx.sa_intv)
    assert np.array_equal(h.bwt_words(), idx.bwt_words())
    assert np.array_equal(h.sa(), idx.sa())
    nb = idx.l_pac // 4 + (1 if idx.l_pac % 4 else 0)
    assert np.array_equal(h.pac()[:nb], idx.pac()[:nb])
    h.close()
