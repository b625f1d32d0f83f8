import pytest
from hypothesis import given, settings, strategies as st


def test_basic_allocation():
    from llama_cpp.kv_pages import OutOfPages, PageAllocator
    a = PageAllocator(8, 16)
    assert a.ensure(0, 1) and a.table(0) == [0]
    assert not a.ensure(0, 16)
    assert a.ensure(0, 17) and len(a.table(0)) == 2
    a.ensure(1, 16 * 6)
    assert a.free_pages == 0
    with pytest.raises(OutOfPages):
        a.ensure(2, 1)
    a.release(0)
    assert a.free_pages == 2 and a.table(0) == []
    a.truncate(1, 20)
    assert len(a.table(1)) == 2 and a.free_pages == 6


@settings(max_examples=200, deadline=None)
@given(st.lists(st.tuples(st.sampled_from(["ensure", "release", "truncate"]), st.integers(0, 5), st.integers(0, 300)), max_size=60))
def test_no_page_is_leaked_or_shared(ops):
    from llama_cpp.kv_pages import OutOfPages, PageAllocator
    a = PageAllocator(12, 32)
    for op, seq, n in ops:
        try:
            if op == "ensure":
                a.ensure(seq, n)
            elif op == "release":
                a.release(seq)
            else:
                a.truncate(seq, n)
        except OutOfPages:
            pass
        owned = [p for s in range(6) for p in a.table(s)]
        assert len(owned) == len(set(owned))
        assert len(owned) + a.free_pages == 12
        assert all(0 <= p < 12 for p in owned)
        for s in range(6):
            pass
