"""Known-answer tests (golden vectors of the reference's unit tests, tests/kat.py) against BOTH CPU checkers."""
import pytest

import kat

CHECKS = [kat.check_morton, kat.check_hilbert, kat.check_sfc_keys_box, kat.check_sort, kat.check_node_counts,
          kat.check_rebalance, kat.check_compute_octree, kat.check_octree, kat.check_find_halos, kat.check_span_sfc_range]


@pytest.mark.parametrize("check", CHECKS, ids=[c.__name__ for c in CHECKS])
def test_port(check, port):
    check(port)


@pytest.mark.parametrize("check", CHECKS, ids=[c.__name__ for c in CHECKS])
def test_ref(check, ref):
    check(ref)


def test_port_neighbors_vs_all2all(port):
    kat.check_find_neighbors(port, n=1200)


def test_ref_neighbors_vs_all2all(ref):
    kat.check_find_neighbors(ref, n=600)
