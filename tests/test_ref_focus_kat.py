"""The reference build (oracle/_ref) against the reference's own known answers for the focus-tree functions: pins
the checker that tests/test_gpu_focus*.py compare the CUDA kernels with."""
import focus_kat


class RefImpl:
    def __init__(self, ref):
        self.ref = ref

    def enforce(self, keys, oc, ops):
        return self.ref.enforce_keys(keys, oc, ops)

    def protect(self, oc, ops):
        return self.ref.protect_ancestors(oc, ops)

    def decision(self, oc, counts, macs, fs, fe, bucket):
        return self.ref.rebalance_decision_essential(oc, counts, macs, fs, fe, bucket)


def test_key_enforcement(ref):
    focus_kat.check_key_enforcement(RefImpl(ref), ref.build_octree)


def test_protect_ancestors(ref):
    focus_kat.check_protect_ancestors(RefImpl(ref), ref.build_octree)


def test_rebalance_decision(ref):
    focus_kat.check_rebalance_decision(RefImpl(ref), ref.build_octree, ref.upsweep_sum)
