"""Host-side helpers that need no GPU: pinned-staging NUMA parsing, the strategy file the BERT
workload imports, and the headers' new entry points being exported by the library."""
import importlib
import json
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]


def test_cpulist_parsing(ffk):
    h = importlib.import_module("flexflow_train_b200.hostmem")
    assert h._parse_cpulist("0-3,8,10-11\n") == [0, 1, 2, 3, 8, 10, 11]
    assert h._parse_cpulist("") == []
    info = h.bind_host_thread_to_gpu_node(0)   # no GPU here: must report, not raise
    assert info["bound"] in (False, True) and "gpu_numa_node" in info


def test_strategy_file_names_a_tensor_parallel_degree():
    s = json.loads((ROOT / "strategies" / "bert_tp2.json").read_text())
    assert int(s["tensor_parallel"]) == 2


def test_peer_memory_and_static_entry_points_are_exported(ffk):
    lib = ffk.lib()
    for name in ("ffk_p2p_create", "ffk_p2p_connect", "ffk_p2p_local_buffer", "ffk_p2p_all_reduce_sum",
                 "ffk_p2p_status", "ffk_p2p_destroy", "ffk_layer_norm_backward_static"):
        assert name in lib.protos and hasattr(lib.cdll, name)


def test_movement_model_reproduces_measured_collectives(ffk):
    """The NVLink model of the cost estimator against the all-reduce times measured on B200
    (profiles/r01_summary.md), within 25 %."""
    ce = importlib.import_module("flexflow_train_b200.cost_estimator")
    for nbytes, ranks, measured_ms in ((32768, 8, 0.0149), (33.563e6, 8, 0.1387), (33.563e6, 2, 0.0778)):
        got = ce.estimate_movement_ms(int(nbytes), "all_reduce", ranks)
        assert abs(got - measured_ms) <= 0.25 * measured_ms, (nbytes, ranks, got, measured_ms)
    assert ce.estimate_movement_ms(0, "all_reduce", 8) == 0.0
    assert ce.estimate_movement_ms(1 << 20, "all_reduce", 1) == 0.0
    assert ce.estimate_movement_ms(1 << 30, "copy", 2) > ce.estimate_movement_ms(1 << 20, "copy", 2)


def test_strategy_reader_understands_both_forms(ffk, tmp_path):
    S = importlib.import_module("flexflow_train_b200.strategy")
    assert S.load_strategy(ROOT / "strategies" / "bert_tp2.json").tensor_parallel == 2
    # a parallel computation graph in the reference's v1 layout: input -> Repartition(sample dim, 4)
    # -> Replicate(2) -> Linear -> Reduction(2) -> Combine(sample dim, 4)
    def layer(kind, value):
        return {"op_attrs": {"type": kind, "value": value}, "name": None}
    pcg = {"raw_graph": {
        "node_labels": {"0": layer("InputAttrs", {}),
                        "1": layer("RepartitionAttrs", {"repartition_dim": {"value": 0}, "repartition_degree": 4}),
                        "2": layer("ReplicateAttrs", {"replicate_degree": 2}),
                        "3": layer("LinearAttrs", {"out_channels": 3072, "use_bias": True}),
                        "4": layer("ReductionAttrs", {"reduction_degree": 2}),
                        "5": layer("CombineAttrs", {"combine_dim": 0, "combine_degree": 4})},
        "output_labels": {},
        "graph": {"nodes": [0, 1, 2, 3, 4, 5],
                  "edges": [{"srcNode": i, "srcIdx": 0, "dstNode": i + 1, "dstIdx": 0} for i in range(5)]}}}
    p = tmp_path / "pcg.json"
    p.write_text(json.dumps(pcg))
    s = S.load_strategy(p)
    assert s.tensor_parallel == 2 and s.data_parallel == 4
    assert sorted(f["op"] for f in s.parallel_ops) == ["combine", "reduction", "repartition", "replicate"]


def _v1(labels, edges):
    """A V1ParallelComputationGraph document the way nlohmann writes it: the int-keyed maps as
    [key, value] pairs, variants as {"type": key, "value": record}."""
    return {"raw_graph": {
        "node_labels": [[i, {"op_attrs": {"type": k, "value": v}, "name": None}] for i, (k, v) in enumerate(labels)],
        "output_labels": [],
        "graph": {"nodes": list(range(len(labels))),
                  "edges": [{"srcNode": a, "srcIdx": 0, "dstNode": b, "dstIdx": i} for a, b, i in edges]}}}


def test_weight_replication_is_not_tensor_parallelism(ffk):
    """What ParallelComputationGraphBuilder really emits for a data-parallel Linear of degree 8
    (parallel_computation_graph_builder.cc:588-634, generate_weight_transform): the input is
    repartitioned along the batch, and EVERY weight gets a Replicate{8} behind its Weight node.
    That graph is data parallel 8, tensor parallel 1."""
    S = importlib.import_module("flexflow_train_b200.strategy")
    labels = [("input", {}),
              ("partition_distributed", {"repartition_dim": {"value": 0}, "repartition_degree": 8}),
              ("weight", {"tensor_shape": {}}), ("replicate_distributed", {"replicate_degree": 8}),
              ("weight", {"tensor_shape": {}}), ("replicate_distributed", {"replicate_degree": 8}),
              ("linear", {"out_channels": 3072, "use_bias": True}),
              ("combine_distributed", {"combine_dim": {"value": 0}, "combine_degree": 8})]
    edges = [(0, 1, 0), (2, 3, 0), (4, 5, 0), (1, 6, 0), (3, 6, 1), (5, 6, 2), (6, 7, 0)]
    s = S.parse_strategy(_v1(labels, edges), world_size=8)
    assert (s.tensor_parallel, s.data_parallel, s.weight_replication) == (1, 8, 8)
    lin = [p for p in s.placements if p.kind == "linear"][0]
    assert [o.op for o in lin.inputs[0]] == ["repartition"] and lin.inputs[0][0].dim == 0
    assert [[o.op for o in w] for w in lin.weights] == [["replicate"], ["replicate"]]
    assert [o.op for o in lin.output] == ["combine"]
    # weight replication alone (no activation op written) still reads as data parallelism
    s2 = S.parse_strategy(_v1(labels[:1] + labels[2:7], [(1, 2, 0), (3, 4, 0), (0, 5, 0), (2, 5, 1), (4, 5, 2)]))
    assert (s2.tensor_parallel, s2.data_parallel) == (1, 8)


def test_tensor_parallel_block_is_placed_per_layer(ffk):
    """Megatron-style feed-forward block in the reference's parallel ops: Replicate(2) in front of a
    column-parallel Linear (weight sharded along out-channels), a row-parallel Linear (weight
    sharded along in-channels) and the Reduction(2) behind it, inside a batch split of 4; the
    weights also carry the data-parallel Replicate(4)."""
    import pytest
    S = importlib.import_module("flexflow_train_b200.strategy")
    labels = [("input", {}),                                                                   # 0
              ("partition_distributed", {"repartition_dim": {"value": 0}, "repartition_degree": 4}),  # 1
              ("replicate_distributed", {"replicate_degree": 2}),                                  # 2
              ("weight", {}), ("partition_distributed", {"repartition_dim": 0, "repartition_degree": 2}),  # 3 4
              ("replicate_distributed", {"replicate_degree": 4}),                                  # 5
              ("linear", {"out_channels": 3072, "use_bias": False}),                               # 6
              ("weight", {}), ("partition_distributed", {"repartition_dim": 1, "repartition_degree": 2}),  # 7 8
              ("replicate_distributed", {"replicate_degree": 4}),                                  # 9
              ("linear", {"out_channels": 768, "use_bias": False}),                                # 10
              ("reduce_distributed", {"reduction_degree": 2}),                                     # 11
              ("layer_norm", {"elementwise_affine": True})]                                        # 12
    edges = [(0, 1, 0), (1, 2, 0), (3, 4, 0), (4, 5, 0), (2, 6, 0), (5, 6, 1),
             (7, 8, 0), (8, 9, 0), (6, 10, 0), (9, 10, 1), (10, 11, 0), (11, 12, 0)]
    doc = _v1(labels, edges)
    s = S.parse_strategy(doc, world_size=8)
    assert (s.tensor_parallel, s.data_parallel) == (2, 4)
    kinds = [p.kind for p in s.placements]
    assert kinds == ["input", "linear", "linear", "layer_norm"]
    col, row = s.placements[1], s.placements[2]
    assert [(o.op, o.degree) for o in col.inputs[0]] == [("repartition", 4), ("replicate", 2)]
    assert [(o.op, o.dim, o.degree) for o in col.weights[0]] == [("repartition", 0, 2), ("replicate", 0, 4)]
    assert [(o.op, o.dim) for o in row.weights[0]][0] == ("repartition", 1)
    assert [(o.op, o.degree) for o in row.output] == [("reduction", 2)]
    assert row.producers == [col.node]
    with pytest.raises(ValueError):
        S.parse_strategy(doc, world_size=6)       # tp 2 x dp 4 != 6
    # the variant may also be written by index
    doc["raw_graph"]["node_labels"][11][1]["op_attrs"] = {"index": S.PCG_OPERATOR_KEYS.index("reduce_distributed"),
                                                         "value": {"reduction_degree": 2}}
    assert S.parse_strategy(doc).tensor_parallel == 2


def test_exported_bert_graph_reads_back_with_its_placement(ffk, tmp_path):
    """pcg_export.bert_pcg writes the BERT graph of lib/models in the reference's v1 layout with a
    tensor-parallel placement; strategy.load_strategy must recover the degrees and, layer by layer,
    where the parallel operators sit."""
    S = importlib.import_module("flexflow_train_b200.strategy")
    X = importlib.import_module("flexflow_train_b200.pcg_export")
    doc = X.bert_pcg(num_layers=3, hidden=64, heads=4, dim_ff=128, seq=16, batch=32, vocab=208, tensor_parallel=2,
                     data_parallel=4)
    p = tmp_path / "bert_tp2_dp4.v1.json"
    X.write(p, doc)
    st = S.load_strategy(p, world_size=8)
    assert (st.tensor_parallel, st.data_parallel) == (2, 4)
    kinds = [pl.kind for pl in st.placements]
    assert kinds.count("multi_head_attention") == 3 and kinds.count("linear") == 7 and kinds.count("layer_norm") == 6
    att = next(pl for pl in st.placements if pl.kind == "multi_head_attention")
    assert [(o.op, o.degree) for o in att.inputs[0]] == [("repartition", 4), ("replicate", 2)]
    assert [(o.op, o.dim, o.degree) for o in att.weights[0]] == [("repartition", 1, 2), ("replicate", 0, 4)]
    assert [(o.op, o.degree) for o in att.output] == [("reduction", 2)]
    lins = [pl for pl in st.placements if pl.kind == "linear"]
    assert [(o.op, o.dim) for o in lins[0].weights[0]][0] == ("repartition", 1)      # column-parallel
    assert [(o.op, o.dim) for o in lins[1].weights[0]][0] == ("repartition", 0)      # row-parallel
    assert [(o.op, o.degree) for o in lins[1].output] == [("reduction", 2)]
    assert [(o.op, o.dim, o.degree) for o in lins[-1].output] == [("combine", 2, 2)]  # vocabulary-parallel head
    ln = next(pl for pl in st.placements if pl.kind == "layer_norm")
    assert [[(o.op, o.degree) for o in w] for w in ln.weights] == [[("replicate", 8)], [("replicate", 8)]]
    # the committed strategy files are such graphs
    for name, world in (("bert_base_tp2_dp1.v1.json", 2), ("bert_base_tp2_dp4.v1.json", 8)):
        st = S.load_strategy(ROOT / "strategies" / name, world_size=world)
        assert st.tensor_parallel == 2 and len([pl for pl in st.placements if pl.kind == "multi_head_attention"]) == 12
    # pure data parallelism exported the same way is not mistaken for tensor parallelism
    st = S.parse_strategy(X.bert_pcg(num_layers=1, hidden=64, heads=4, dim_ff=128, seq=16, batch=32, vocab=208,
                                     tensor_parallel=1, data_parallel=8), world_size=8)
    assert (st.tensor_parallel, st.data_parallel) == (1, 8)


def _kcut(ffk, M, N, K, pairs):
    import ctypes

    L = ffk.lib()
    plan = (ctypes.c_int * 5)()
    taken = L.call("ffk_gemm_kcut_plan", M, N, K, pairs, plan)
    segs = []
    for pair in range(pairs):
        buf = (ctypes.c_int * (6 * 1024))()
        n = L.call("ffk_gemm_kcut_segments", M, N, K, pairs, pair, buf, 1024)
        assert n <= 1024
        segs.append([tuple(buf[6 * i + j] for j in range(6)) for i in range(n)])
    return taken, tuple(plan), segs


def test_gemm_k_cut_schedule_covers_every_k_block_once(ffk):
    """The segment lists the CTA pairs walk (the kernel and this test call the same code):
    every (tile, k-block) exactly once; each parked partial has its own slot, below the pair
    count (the scratch holds one slot per pair); the slots a tile's owner reads are exactly the
    slots parked for that tile; a pair never parks after it has waited (so waits cannot chain);
    pairs that share a tile row / column start their pieces at the same k-block (one read from
    HBM serves them); and the longest pair is within a few k-blocks of the ideal share."""
    import random

    rng = random.Random(5)
    shapes = [(4096, 4096, 8192, 74), (4096, 1024, 32768, 74), (768, 768, 32768, 74), (3072, 768, 32768, 74),
              (512, 512, 4096, 74), (4096, 512, 512, 74), (1000, 1304, 4104, 74), (8192, 4096, 4096, 74)]
    shapes += [(rng.randint(129, 6000), rng.randint(129, 6000), rng.randint(8, 40000), rng.choice([74, 66, 10, 2, 128]))
               for _ in range(80)]
    cut_seen = 0
    for M, N, K, pairs in shapes:
        taken, plan, segs = _kcut(ffk, M, N, K, pairs)
        whole, cut, pieces, piece_k, tails = plan
        tiles, num_k = ((M + 255) // 256) * ((N + 255) // 256), (K + 63) // 64
        cover, parked, owned, slots = set(), {}, {}, set()
        for pair, lst in enumerate(segs):
            waited = False
            for (t, a, b, role, slot, count) in lst:
                assert 0 <= t < tiles and 0 <= a < b <= num_k, (M, N, K, pairs, t, a, b)
                for kb in range(a, b):
                    assert (t, kb) not in cover
                    cover.add((t, kb))
                if role == 0:
                    assert (a, b) == (0, num_k)
                elif role == 1:
                    assert a > 0 and not waited, "a pair parks after it has waited"
                    assert 0 <= slot < pairs and slot not in slots
                    slots.add(slot)
                    parked.setdefault(t, []).append(slot)
                else:
                    assert role == 2 and a == 0 and b < num_k and count >= 1
                    owned[t] = list(range(slot, slot + count))
                    waited = True
        assert len(cover) == tiles * num_k
        assert {t: sorted(v) for t, v in parked.items()} == owned
        if not taken:
            assert plan == (0, 0, 0, 0, 0) and not parked
            continue
        cut_seen += 1
        assert whole % pairs == 0 and whole + cut == tiles and 0 < cut < pairs
        # cut points are the same for every cut tile
        starts = {a for lst in segs for (t, a, b, role, _, _) in lst if role != 0}
        assert starts == ({j * piece_k for j in range(pieces)} if pieces else {0, piece_k})
        loads = [sum(b - a for (_, a, b, _, _, _) in lst) for lst in segs]
        uncut = -(-tiles // pairs) * num_k
        assert max(loads) < uncut, (M, N, K, pairs, max(loads), uncut)
    assert cut_seen >= 25
    # the shapes the schedule exists for, and one it must leave alone
    assert _kcut(ffk, 4096, 1024, 32768, 74)[0] == 0                             # encoder weight gradient, 64 tiles on 74 pairs: measured slower when cut
    assert _kcut(ffk, 2000, 1304, 1032, 74)[:2] == (1, (0, 48, 0, 12, 2))        # 48 tiles: heads of 12 k-blocks, two tails per helper
    assert _kcut(ffk, 4096, 4096, 8192, 74)[:2] == (1, (222, 34, 2, 64, 0))      # configs[1] weight gradient: 3.46 waves
    assert _kcut(ffk, 768, 768, 32768, 74)[:2] == (1, (0, 9, 8, 64, 0))          # BERT-base weight gradient: 9 tiles
    assert _kcut(ffk, 8192, 4096, 4096, 74)[0] == 0                              # 512 tiles = 6.92 waves


def test_peer_grad_sync_slices_tile_every_bucket_once(ffk, monkeypatch):
    """PeerGradSync's bookkeeping (host logic of the N > 1 path) with the device taken away: a
    stand-in peer group records the collectives. Over one backward pass (layers in reverse) plus
    finish(), the slices sent for each bucket must tile [0, bucket) exactly once, from the tail
    down, on 16-byte boundaries, each at least min_bytes long except the remainder; `summed(g)` is
    the view of the sum region at g's offset; the K-cut GEMM schedule is off while the object lives."""
    import torch

    P = importlib.import_module("flexflow_train_b200.parallel_ops")
    calls, switches = [], []

    class FakeK:
        @staticmethod
        def set_gemm_kcut(on):
            switches.append(bool(on))

    class FakePeer:
        K = FakeK

        def __init__(self, nbytes, group=None):
            self.world, self.rank = 2, 0
            self.nbytes = (int(nbytes) + 255) // 256 * 256
            self.bytes = torch.zeros(self.nbytes, dtype=torch.uint8)

        def tensor(self, off, shape, dtype):
            n = 1
            for s in shape:
                n *= s
            nb = n * torch.empty(0, dtype=dtype).element_size()
            assert off % 16 == 0 and off + nb <= self.nbytes
            return self.bytes[off:off + nb].view(dtype).view(tuple(shape))

        def all_reduce_sum(self, in_off, out_off, count, dtype, stream=None, background_blocks=0):
            calls.append((in_off, out_off, count, dtype))

        def check(self):
            pass

        def close(self):
            pass

    class Dummy:
        def __init__(self, *a, **k):
            pass

        def record(self, *a):
            pass

        def wait_event(self, *a):
            pass

    monkeypatch.setattr(P, "PeerMemoryGroup", FakePeer)
    monkeypatch.setattr(torch.cuda, "Stream", Dummy)
    monkeypatch.setattr(torch.cuda, "Event", Dummy)
    monkeypatch.setattr(torch.cuda, "current_stream", lambda *a: Dummy())

    class Layer:
        def __init__(self, grads):
            self.grads = grads

    class Model:
        """Two buckets (bf16 weights, fp32 norms), gradients laid out layer by layer like training.Model."""

        def __init__(self):
            sizes16 = [4096 * 300, 1000, 4096 * 300, 7, 4096 * 500]
            sizes32 = [1024, 1024, 3, 1024]
            self.spec = [(torch.bfloat16, sizes16), (torch.float32, sizes32)]
            self.layers = [Layer([]) for _ in range(5)]
            self._lay([torch.zeros(sum((n + 7) // 8 * 8 for n in sizes16), dtype=torch.bfloat16),
                       torch.zeros(sum((n + 3) // 4 * 4 for n in sizes32), dtype=torch.float32)])

        def _lay(self, flats):
            self.grad_buckets = flats
            for l in self.layers:
                l.grads = []
            for flat, (dt, sizes) in zip(flats, self.spec):
                align, off = 16 // flat.element_size(), 0
                for i, n in enumerate(sizes):
                    self.layers[i % len(self.layers)].grads.append(flat[off:off + n])
                    off += (n + align - 1) // align * align
            self.grads = [g for l in self.layers for g in l.grads]

        def rehome_gradients(self, flats):
            self._lay(flats)

    model = Model()
    min_bytes = 1 << 20
    sync = P.PeerGradSync(model, min_bytes=min_bytes, background_blocks=0)
    assert switches == [False]
    for step in range(2):                       # the bookkeeping must reset for the next step
        calls.clear()
        for l in reversed(model.layers):
            sync.after_layer(l)
        sync.finish()
        for bi, (goff, soff, dt, es, numel) in enumerate(sync.slots):
            mine = [(i - goff, o - soff, c) for (i, o, c, d) in calls if d == dt]
            assert mine and all(i == o and i % 16 == 0 for i, o, _ in mine)
            spans = [(i, i + c * es) for i, _, c in mine]
            assert spans[0][1] == numel * es and spans[-1][0] == 0
            assert all(a[0] == b[1] for a, b in zip(spans, spans[1:])), "slices must follow each other downwards"
            assert all(hi - lo >= min_bytes for lo, hi in spans[:-1])
    for g in model.grads:
        sm = sync.summed(g)
        bi = next(i for i, f in enumerate(model.grad_buckets)
                  if f.dtype == g.dtype and f.data_ptr() <= g.data_ptr() < f.data_ptr() + f.numel() * f.element_size())
        flat, smf = model.grad_buckets[bi], sync.sums[bi]
        assert sm.shape == g.shape and sm.data_ptr() - smf.data_ptr() == g.data_ptr() - flat.data_ptr()
    sync.close()
    assert switches == [False, True]
