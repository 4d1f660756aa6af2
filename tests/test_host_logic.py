"""Host-side logic that needs neither a GPU nor a process group."""
import inspect

import pytest
import torch

from distributed_pipeline_b200 import FlatLayout, FlatState, TrainLoop, dist_util, logger, update_ema
from distributed_pipeline_b200.engine import Engine, ema_update
from distributed_pipeline_b200.optim import FusedAdamW
from distributed_pipeline_b200.synthetic import SyntheticDiffuSeq, diffuseq_shapes, numel


def test_layout_alignment_and_padding():
    lay = FlatLayout.from_numels([5, 0, 130, 7, 32], 4)
    assert lay.offsets == (0, 32, 32, 192, 224) and lay.n_elems == 256
    assert all(o * 4 % 128 == 0 for o in lay.offsets)
    lay2 = FlatLayout.from_numels([5, 0, 130, 7, 32], 2)
    assert all(o * 2 % 128 == 0 for o in lay2.offsets) and lay2.n_elems % 64 == 0
    assert lay.n_params == 174
    assert lay.segment_of(32, 192) == [1, 2, 3]
    with pytest.raises(ValueError):
        FlatLayout.from_numels([1], 4, align_bytes=24)
    full = FlatLayout.from_numels([numel(s) for s in diffuseq_shapes()], 4)
    assert full.n_params == 91225274 and full.n_elems - full.n_params < 209 * 32


def test_flat_state_repoints_parameters_and_keeps_grad_views():
    torch.manual_seed(0)
    model = SyntheticDiffuSeq(n_layers=1, d=64, ffn=128, vocab=50, emb=16, t_hidden=32, max_pos=16)
    before = [p.detach().clone() for p in model.parameters()]
    st = FlatState(list(model.parameters()), n_ema=2)
    for p, b, pv, gv in zip(model.parameters(), before, st.param_views, st.grad_views):
        assert torch.equal(p, b) and p.data_ptr() == pv.data_ptr() and p.grad.data_ptr() == gv.data_ptr()
    for e in st.emas:
        assert torch.equal(e, st.params)  # EMA copies start as copies (reference :111-113)
    # autograd accumulates into the flat gradient buffer, across microbatches
    ids, t = torch.randint(0, 50, (2, 8)), torch.randint(0, 100, (2,))
    model(ids, t).pow(2).mean().backward()
    g1 = st.grads.clone()
    assert g1.abs().sum() > 0
    model(ids, t).pow(2).mean().backward()
    assert torch.allclose(st.grads, 2 * g1, rtol=1e-5, atol=1e-8)
    st.recheck_views()
    # padding between tensors never receives anything
    mask = torch.ones(st.layout.n_elems, dtype=torch.bool)
    for o, n in zip(st.layout.offsets, st.layout.numels):
        mask[o:o + n] = False
    assert not st.grads[mask].any() and not st.params[mask].any()
    next(model.parameters()).grad = None
    with pytest.raises(RuntimeError):
        st.recheck_views()
    with pytest.raises(TypeError):
        FlatState([torch.nn.Parameter(torch.zeros(3)), torch.nn.Parameter(torch.zeros(3, dtype=torch.float64))], 0)


def test_no_cpu_fallback_anywhere():
    lay = FlatLayout.from_numels([10], 4)
    with pytest.raises(RuntimeError, match="no CPU implementation"):
        Engine(lay, torch.float32, "cpu")
    with pytest.raises(RuntimeError, match="no CPU implementation"):
        ema_update(torch.zeros(4), torch.zeros(4), 0.9)
    with pytest.raises(RuntimeError):
        update_ema([torch.zeros(4)], [torch.zeros(4)], 0.9)
    with pytest.raises(RuntimeError):
        FusedAdamW([torch.nn.Parameter(torch.zeros(4))])
    with pytest.raises(RuntimeError, match="CUDA"):
        TrainLoop(model=torch.nn.Linear(2, 2), data=iter([]), batch_size=1, microbatch=1, lr=1e-3, ema_rate="0.9",
                  log_interval=1, save_interval=1, resume_checkpoint="")


def test_trainloop_signature_matches_reference_surface():
    """Constructor keywords (reference utils/trainer.py:45-65) in the same order,
    same defaults; every reference method present."""
    sig = inspect.signature(TrainLoop.__init__)
    names = [n for n in sig.parameters if n != "self"]
    ref = ["model", "data", "batch_size", "microbatch", "lr", "ema_rate", "log_interval", "save_interval",
           "resume_checkpoint", "weight_decay", "learning_steps", "checkpoint_path", "gradient_clipping",
           "eval_data", "eval_interval", "eval_callbacks"]
    assert names[:len(ref)] == ref
    assert all(sig.parameters[n].kind is inspect.Parameter.KEYWORD_ONLY for n in names)
    defaults = {n: sig.parameters[n].default for n in ref[9:]}
    assert defaults == dict(weight_decay=0.0, learning_steps=0, checkpoint_path='', gradient_clipping=-1.,
                            eval_data=None, eval_interval=-1, eval_callbacks=())
    for extra in names[len(ref):]:
        assert sig.parameters[extra].default is not inspect.Parameter.empty
    for meth in ["log_loss_dict", "compute_losses", "backward_from_losses", "get_batch_length", "run_loop",
                 "run_step", "_zero_grad", "_common_forward", "forward_only", "forward_backward", "optimize",
                 "grad_clip", "_anneal_lr", "_log_grad_norm", "log_step", "save", "_save_checkpoint", "_save_opt",
                 "_master_params_to_state_dict", "_state_dict_to_master_params", "parse_resume_step_from_filename",
                 "find_resume_checkpoint", "find_ema_checkpoint", "find_opt_checkpoint", "_load_and_sync_parameters",
                 "_load_ema_parameters", "_load_optimizer_state", "__call__"]:
        assert hasattr(TrainLoop, meth), meth
    assert isinstance(inspect.getattr_static(TrainLoop, "backward_from_losses"), staticmethod)
    assert TrainLoop.get_batch_length({"a": torch.zeros(5, 2)}) == 5
    assert TrainLoop.get_batch_length([torch.zeros(3)]) == 3
    with pytest.raises(TypeError):
        TrainLoop.get_batch_length("x")
    assert TrainLoop.parse_resume_step_from_filename("/x/model_000123.pt") == 123
    assert list(inspect.signature(update_ema).parameters) == ["target_params", "source_params", "rate"]
    assert inspect.signature(update_ema).parameters["rate"].default == 0.99


def test_dist_util_single_process_degrades_like_reference(monkeypatch):
    assert dist_util.get_rank() == 0 and dist_util.get_world_size() == 1
    assert dist_util.barrier() is None
    dist_util.sync_params([torch.zeros(3)])  # no-op without a group
    assert dist_util.dev().type in ("cpu", "cuda")
    port = dist_util.find_free_port()
    assert 0 < port < 65536
    for name in ["setup_dist", "is_available", "is_initialized", "get_rank", "get_world_size", "barrier", "dev",
                 "load_state_dict", "broadcast", "sync_params", "find_free_port"]:
        assert callable(getattr(dist_util, name))


def test_logger_running_mean_and_dump(tmp_path):
    logger.configure(dir=str(tmp_path))
    for v in (1.0, 2.0, 6.0):
        logger.logkv_mean("grad_norm", v)
    logger.logkv("step", 7)
    out = logger.dumpkvs()
    assert out["grad_norm"] == pytest.approx(3.0) and out["step"] == 7
    assert logger.getkvs() == {}
    assert (tmp_path / "progress.csv").exists()
    assert logger.get_current().dir == str(tmp_path)


class _StubEngine:
    """Engine stand-in for host-only tests of FusedAdamW's bookkeeping (no kernels run)."""

    def __init__(self, layout, dtype, device, ema_rates=(), betas=(0.9, 0.999), eps=1e-8, **kw):
        self.layout, self.betas, self.eps, self.world_size, self.nvls = layout, tuple(betas), eps, 1, False
        self.calls = []

    def bind(self, *a):
        pass

    def fused_update(self, lr, wd, step, **kw):
        self.calls.append(("update", lr, wd, step, kw))

    def grad_norm(self, inv_scale, max_norm):
        self.calls.append(("norm", inv_scale, max_norm))

    def zero_grads(self):
        self.calls.append(("zero",))


def test_fused_adamw_state_dict_is_torch_adamw_layout(monkeypatch):
    """opt_NNNNNN.pt must be interchangeable with torch.optim.AdamW's (reference :165-173, :296-302)."""
    from distributed_pipeline_b200 import optim as optim_mod
    monkeypatch.setattr(optim_mod, "Engine", _StubEngine)
    torch.manual_seed(0)
    params = [torch.nn.Parameter(torch.randn(5, 3)), torch.nn.Parameter(torch.randn(7)), torch.nn.Parameter(torch.randn(2, 2, 2))]
    ours = optim_mod.FusedAdamW(params, lr=1e-3, weight_decay=0.01, ema_rates=[0.99])
    assert ours.state_dict()["state"] == {} and ours.state == {}        # like torch before the first step
    ours.clip_grad_norm(1.0)
    ours.step()
    assert [c[0] for c in ours.engine.calls] == ["norm", "update"]
    assert ours.engine.calls[1][3] == 1 and ours.engine.calls[1][4]["use_clip_coef"] is True
    ours.step()
    assert ours.engine.calls[2][4]["use_clip_coef"] is False              # the coefficient is used once
    assert ours.ema_done == [True]

    stock_params = [torch.nn.Parameter(p.detach().clone()) for p in params]
    stock = torch.optim.AdamW(stock_params, lr=1e-3, weight_decay=0.01)
    for p in stock_params:
        p.grad = torch.randn_like(p)
    stock.step()
    stock.step()
    a, b = ours.state_dict(), stock.state_dict()
    assert set(a) == set(b) == {"state", "param_groups"}
    assert list(a["state"]) == list(b["state"]) == [0, 1, 2]
    for i in range(3):
        assert set(a["state"][i]) == set(b["state"][i]) == {"step", "exp_avg", "exp_avg_sq"}
        assert float(a["state"][i]["step"]) == float(b["state"][i]["step"]) == 2.0
        assert a["state"][i]["exp_avg"].shape == b["state"][i]["exp_avg"].shape
    assert a["param_groups"][0]["params"] == b["param_groups"][0]["params"] == [0, 1, 2]
    assert set(b["param_groups"][0]) <= set(a["param_groups"][0])
    # a stock checkpoint loads into the fused optimizer (moments land in the flat buffers) ...
    ours.load_state_dict(b)
    assert ours.step_count == 2
    for i, p in enumerate(stock_params):
        assert torch.equal(ours.flat.exp_avg_views[i], stock.state[p]["exp_avg"])
        assert torch.equal(ours.flat.exp_avg_sq_views[i], stock.state[p]["exp_avg_sq"])
    # ... and ours loads into a stock AdamW
    fresh = torch.optim.AdamW([torch.nn.Parameter(p.detach().clone()) for p in params], lr=1e-3, weight_decay=0.01)
    fresh.load_state_dict(ours.state_dict())
    assert float(next(iter(fresh.state.values()))["step"]) == 2.0
    # mismatching layouts are refused
    with pytest.raises(ValueError):
        ours.load_state_dict({"state": {}, "param_groups": [{"params": [0]}]})
    with pytest.raises(ValueError):
        ours.zero_grad(set_to_none=True)
    with pytest.raises(NotImplementedError):
        ours.step(closure=lambda: None)


def test_round2_host_logic(monkeypatch):
    """Layout / reducer / logger / dist_util details fixed in round 2, all without a GPU."""
    # a trailing zero-element tensor sits at offset == n_elems: it belongs to the range that ends there
    lay = FlatLayout.from_numels([100, 7, 0], 4)
    assert lay.offsets[2] == lay.n_elems
    assert lay.segment_of(0, lay.n_elems) == [0, 1, 2]
    assert lay.segment_of(0, lay.offsets[1]) == [0]
    # running mean of n equal samples in one update == n single updates
    a, b = logger.Logger(None), logger.Logger(None)
    a.logkv_mean("x", 1.0)
    b.logkv_mean("x", 1.0)
    for _ in range(5):
        a.logkv_mean("x", 3.5)
    b.logkv_mean_n("x", 3.5, 5)
    assert abs(a.name2val["x"] - b.name2val["x"]) < 1e-12 and a.name2cnt["x"] == b.name2cnt["x"] == 6
    # LOCAL_RANK alone decides, as in the reference (basic_utils/dist_util.py:66): with the other env://
    # variables missing init_process_group fails and setup_dist falls back to a single process
    monkeypatch.setenv("LOCAL_RANK", "0")
    for k in ("RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT"):
        monkeypatch.delenv(k, raising=False)
    monkeypatch.setitem(dist_util._state, "setup_done", None)
    assert dist_util.setup_dist(silent=True) is False
    assert dist_util.get_world_size() == 1
    # the new constructor keywords are optional and keyword-only
    sig = inspect.signature(TrainLoop.__init__)
    for k in ("schedule_sampler", "diffusion_steps", "timestep_key", "use_fp16", "grads_after_step", "grad_allreduce"):
        assert sig.parameters[k].kind is inspect.Parameter.KEYWORD_ONLY
    assert sig.parameters["schedule_sampler"].default is None


def test_reducer_waits_only_for_parameters_that_get_gradients():
    """A frozen tensor never fires a hook: it must not hold its bucket (and every later one) back."""
    from distributed_pipeline_b200.reducer import BucketReducer

    class FakeFlat:
        pass

    class FakeEngine:
        world_size = 2
        buckets = [(300, 100), (0, 300)]          # tail first

    ps = [torch.nn.Parameter(torch.zeros(100)), torch.nn.Parameter(torch.zeros(200), requires_grad=False),
          torch.nn.Parameter(torch.zeros(100))]
    flat = FakeFlat()
    flat.param_list = ps
    flat.layout = FlatLayout((0, 100, 300), (100, 200, 100), 400, 4)
    r = BucketReducer(flat, FakeEngine())
    assert r.bucket_params == [[2], [0, 1]]
    assert r._hooked == [1, 1] and r._pending == [1, 1] and r._ready == [False, False]
    assert len(r._handles) == 2                   # no hook on the frozen tensor
    r.remove()
