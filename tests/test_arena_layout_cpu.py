"""Properties of the flat-arena layout (engine/variables.py: finalize) that the peer-memory collectives rely on, checked
on the CPU box for random variable sets, both placements and 1-8 owners: the CUDA-only steps (tap-transpose table,
device allocation) are stubbed out, the layout code itself is the one that runs on the GPU."""
import torch
from hypothesis import given, settings, strategies as st

from distributed_tensorflow_framework_b200.engine import variables as V
from distributed_tensorflow_framework_b200.engine.variables import ConstantInitializer, XavierInitializer


class _FakeCuda(object):
    type = "cuda"


KINDS = ["conv_w", "fc_w", "deconv_w", "bias", "generic"]
var_spec = st.tuples(st.sampled_from(KINDS), st.integers(1, 5), st.integers(1, 70), st.integers(1, 40))


def _make_store(specs):
    store = V.VariableStore(torch.device("cpu"), seed=3)
    for i, (kind, k, ci, co) in enumerate(specs):
        if kind in ("conv_w", "deconv_w"):
            shape = [k, k, ci, co]
        elif kind == "fc_w":
            shape = [ci * k, co]
        else:
            shape = [co]
        init = XavierInitializer() if len(shape) > 1 else ConstantInitializer(0.25 + i)
        store.get_variable("v%d/%s" % (i, kind), shape, init, kind=kind)
    store.get_variable("stats/moving_mean", [7], ConstantInitializer(0.0), trainable=False)
    return store


@settings(max_examples=60, deadline=None)
@given(specs=st.lists(var_spec, min_size=1, max_size=24), world=st.integers(1, 8),
       placement=st.sampled_from(["bytes", "round_robin"]))
def test_arena_layout_properties(specs, world, placement):
    store = _make_store(specs)
    before = {v.name: v.master.detach().clone() for v in store.trainable_variables()}
    store.device = _FakeCuda()
    store._build_tap_table = lambda descs, tiles: setattr(store, "_tap_table", None)
    store.finalize(allocator=lambda tag, n, dt: torch.zeros(n, dtype=dt), placement=placement, world=world)
    tv = store.trainable_variables()
    n = store.n_params
    spans = sorted((v.offset, v.offset + v.numel, v.name) for v in tv)
    # every variable on a 128-byte line of the bf16 arenas, no overlap, everything inside the arena
    assert all(lo % V.ARENA_ALIGN == 0 for lo, _, _ in spans)
    assert all(spans[i][1] <= spans[i + 1][0] for i in range(len(spans) - 1)) and spans[-1][1] <= n
    assert n % V.ARENA_ALIGN == 0 and store.master_arena.numel() == n == store.grad_arena.numel()
    # owner ranges: a monotone partition of [0, n); 16-byte boundaries (the collectives move 16-byte units)
    b = store.owner_bounds
    assert len(b) == world + 1 and b[0] == 0 and b[-1] == n and all(b[i] <= b[i + 1] for i in range(world))
    assert all(x % 4 == 0 for x in b)
    small = lambda v: v.kind not in ("conv_w", "fc_w", "deconv_w")  # noqa: E731
    for v in tv:
        # values survive the move into the arena; master / grad / shadow are views of the same slice
        assert torch.equal(v.master.reshape(-1), before[v.name].reshape(-1))
        assert v.master.data_ptr() == store.master_arena[v.offset:].data_ptr()
        assert v.grad.data_ptr() == store.grad_arena[v.offset:].data_ptr()
        assert torch.equal(v.shadow.detach().float().reshape(-1),
                           before[v.name].to(store.compute_dtype).float().reshape(-1))
        assert 0 <= v.owner < world and b[v.owner] <= v.offset < max(b[v.owner + 1], b[v.owner] + 1)
        if small(v):  # kernels read these in fp32: they lie inside the range their owner broadcasts
            lo, hi = store.f32_ranges[v.owner]
            assert lo <= v.offset and v.offset + v.numel <= hi
    if placement == "round_robin" and world > 1:
        assert all(v.owner == i % world for i, v in enumerate(sorted(tv, key=lambda v: v.index)))
        for o in range(world):  # inside an owner's shard: the fp32-read variables first
            mine = sorted((v for v in tv if v.owner == o), key=lambda v: v.offset)
            flags = [small(v) for v in mine]
            assert flags == sorted(flags, reverse=True)
            assert all(b[o] <= v.offset and v.offset + v.numel <= b[o + 1] for v in mine)
    else:
        ordered = sorted(tv, key=lambda v: v.offset)
        flags = [small(v) for v in ordered]
        assert flags == sorted(flags, reverse=True)  # fp32-read variables in front of the tensor-core operands
        lo, hi = store.f32_ranges[0]
        assert all(not small(v) or v.offset + v.numel <= hi for v in tv) and lo == 0
    # the non-trainable statistics stay outside the arenas
    assert store.vars["stats/moving_mean"].master.numel() == 7


@settings(max_examples=60, deadline=None)
@given(specs=st.lists(var_spec, min_size=1, max_size=24), world=st.integers(1, 8),
       placement=st.sampled_from(["bytes", "round_robin"]), bucket_kib=st.sampled_from([1, 4, 64, 4096]))
def test_gradient_buckets_partition_the_arena(specs, world, placement, bucket_kib):
    from distributed_tensorflow_framework_b200.parallel.comm import plan_buckets
    store = _make_store(specs)
    store.device = _FakeCuda()
    store._build_tap_table = lambda descs, tiles: setattr(store, "_tap_table", None)
    store.finalize(allocator=lambda tag, n, dt: torch.zeros(n, dtype=dt), placement=placement, world=world)
    tv = store.trainable_variables()
    buckets = plan_buckets(tv, store.n_params, bucket_kib << 10)
    assert buckets[0][0] == 0 and buckets[-1][1] == store.n_params
    assert all(buckets[i][1] == buckets[i + 1][0] for i in range(len(buckets) - 1))
    seen = [v.name for _, _, vs in buckets for v in vs]
    assert sorted(seen) == sorted(v.name for v in tv) and len(seen) == len(set(seen))
    per = max((bucket_kib << 10) // 4, 1024)
    for i, (lo, hi, vs) in enumerate(buckets):
        assert vs and all(lo <= v.offset and v.offset + v.numel <= hi for v in vs)
        if i < len(buckets) - 1:  # closed at the FIRST boundary past the threshold
            assert hi - lo >= per and (len(vs) == 1 or vs[-1].offset - lo < per)
