"""CPU: pure-Python host logic of the mirror (graphblas_b200) that needs no device."""


def test_slice_index_lists():
    """Vector slices other than `[:]` become the index list they stand for (graphblas_b200.core._slice_index)."""
    import numpy as np
    from graphblas_b200.core import _slice_index

    assert _slice_index(slice(2, 10, 2), 7).tolist() == [2, 4, 6]
    assert _slice_index(slice(None, 3), 7).tolist() == [0, 1, 2]
    assert _slice_index(slice(-2, None), 7).tolist() == [5, 6]
    assert _slice_index(slice(None, None, -3), 7).tolist() == [6, 3, 0]
    assert _slice_index(slice(5, 2), 7).size == 0
    assert _slice_index(slice(1, 4), 7).dtype == np.uint64
