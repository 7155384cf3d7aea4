"""T2: collaborative optimizer end to end on localhost - epochs advance in lock-step, weighted gradient averaging is
correct, a late joiner downloads state, state_dict layout carries local_epoch."""
import threading
import time

import torch

from dalle_b200.collab import Float16Compression, Optimizer, SizeAdaptiveCompression, Uniform8BitQuantization
from dalle_b200.dht import DHT


def _dht(peers=()):
    return DHT(initial_peers=list(peers), start=True, host_maddrs=["/ip4/127.0.0.1/tcp/0"])


def _make(dht, run_id, bs, target=32, lr=0.1, **kw):
    torch.manual_seed(0)
    model = torch.nn.Linear(8, 1)
    opt = Optimizer(dht=dht, run_id=run_id, target_batch_size=target, batch_size_per_step=bs,
                    optimizer=lambda params: torch.optim.SGD(params, lr=lr), params=list(model.parameters()),
                    scheduler=lambda o: torch.optim.lr_scheduler.LambdaLR(o, lambda e: 1.0),
                    matchmaking_time=1.5, averaging_timeout=15, allreduce_timeout=10, reuse_grad_buffers=True,
                    delay_optimizer_step=True, target_group_size=kw.pop("group", None),
                    tracker_opts=dict(min_refresh_period=0.1, default_refresh_period=0.3, max_refresh_period=0.5), **kw)
    return model, opt


def test_two_peers_advance_in_lockstep_with_weighted_average():
    d0 = _dht()
    d1 = _dht([str(m) for m in d0.get_visible_maddrs()])
    try:
        (m0, o0), (m1, o1) = _make(d0, "lock", bs=8, group=2), _make(d1, "lock", bs=8, group=2)
        w_init = m0.weight.detach().clone()
        stop, errors = threading.Event(), []

        def train(model, opt, gval):
            try:
                while not stop.is_set() and opt.local_epoch < 2:
                    for p in model.parameters():
                        p.grad = torch.full_like(p, gval) if p.grad is None else p.grad.add_(gval)
                    opt.step()
                    time.sleep(0.05)
            except Exception as e:  # pragma: no cover
                errors.append(e)

        ts = [threading.Thread(target=train, args=(m0, o0, 1.0)), threading.Thread(target=train, args=(m1, o1, 3.0))]
        [t.start() for t in ts]
        [t.join(60) for t in ts]
        stop.set()
        assert not errors, errors
        assert o0.local_epoch >= 2 and o1.local_epoch >= 2
        # each epoch: averaged gradient = sample-weighted mean of per-peer mean grads, between 1 and 3
        delta = (w_init - m0.weight.detach()) / 0.1 / o0.local_epoch
        assert 0.99 <= delta.mean().item() <= 3.01
        o0._finish_pending(); o1._finish_pending()
        assert torch.allclose(m0.weight, m1.weight, atol=1e-4), (m0.weight, m1.weight)
        sd = o0.state_dict()
        assert sd["state"]["local_epoch"] == o0.local_epoch
        o0.shutdown(); o1.shutdown()
    finally:
        d0.shutdown(); d1.shutdown()


def test_single_peer_uses_local_gradients_and_late_joiner_loads_state():
    d0 = _dht()
    try:
        m0, o0 = _make(d0, "solo", bs=16, target=32, averager_opts=dict(declare_state_period=0.5))
        for _ in range(6):
            for p in m0.parameters():
                p.grad = torch.ones_like(p) if p.grad is None else p.grad.add_(1.0)
            o0.step()
        assert o0.local_epoch == 3
        assert o0.grad_averager.local_samples_accumulated == 0
        o0.state_averager.declare_state_now()
        d1 = _dht([str(m) for m in d0.get_visible_maddrs()])
        m1, o1 = _make(d1, "solo", bs=16, target=32)
        time.sleep(0.5)
        assert o1.load_state_from_peers()
        assert o1.local_epoch == 3 and torch.allclose(m1.weight, m0.weight)
        o0.shutdown(); o1.shutdown(); d1.shutdown()
    finally:
        d0.shutdown()


def test_compression_choice_matches_reference_threshold():
    c = SizeAdaptiveCompression(threshold=2 ** 16 + 1, less=Float16Compression(), greater_equal=Uniform8BitQuantization())
    assert c.choose(65536).name == "fp16" and c.choose(65537).name == "u8"


def test_grad_scaler_defers_unscale_and_update_to_global_steps():
    """HivemindGradScaler (reference hf_trainer.py:26-28): gradients accumulated over micro-steps keep ONE
    loss scale - unscale_ / update are no-ops between global steps and run once inside the global step."""
    from dalle_b200.collab.optimizer import HivemindGradScaler

    d0 = _dht()
    try:
        model, opt = _make(d0, "scaler", bs=8, target=16, lr=0.5)
        scaler = HivemindGradScaler(device="cpu", init_scale=1024.0, growth_interval=1, enabled=True)
        w0 = model.weight.detach().clone()
        x = torch.ones(4, 8)
        for micro in range(2):            # two micro-steps = one global step (target 16, 8 per step)
            loss = model(x).sum()
            scaler.scale(loss).backward()
            assert scaler.unscale_(opt) is False      # deferred: the accumulated gradients stay scaled by 1024
            if micro == 0:
                assert abs(model.weight.grad.mean().item() - 4.0 * 1024.0) < 1e-3
            scaler.step(opt)
            scaler.update()
            if micro == 0:
                assert scaler.get_scale() == 1024.0 and opt.local_epoch == 0
        assert opt.local_epoch == 1
        # the wrapped SGD saw the UNSCALED mean gradient of the two micro-steps: d(sum)/dw = 4 per element
        assert torch.allclose(w0 - model.weight.detach(), torch.full_like(w0, 0.5 * 4.0), atol=1e-4)
        assert scaler.get_scale() == 2048.0           # one update, at the global step
        opt.shutdown()
    finally:
        d0.shutdown()


def test_nan_in_parameters_rolls_back_to_the_local_backup(tmp_path):
    """callback.on_step_end: a non-finite parameter restores model + optimizer + scheduler + epoch from state_path
    (reference callback.py:50-55, 115-127); without a backup it raises."""
    import types

    import pytest

    import callback as cbmod

    d0 = _dht()
    try:
        model, opt = _make(d0, "nanroll", bs=16, target=16, lr=0.1)
        task = types.SimpleNamespace(dht=d0, collaborative_optimizer=opt, model=model, local_public_key=b"k")
        args = types.SimpleNamespace(statistics_expiration=60, backup_every_steps=1, state_path=str(tmp_path / "s.zip"))
        cb = cbmod.CollaborativeCallback(task, args)

        def micro_step():
            for p in model.parameters():
                p.grad = torch.ones_like(p) if p.grad is None else p.grad.add_(1.0)
            opt.step()
            cb.on_step_end(loss=1.0)

        with torch.no_grad():
            model.weight[0, 0] = float("nan")
        with pytest.raises(RuntimeError, match="no backup"):
            cb.on_step_end(loss=1.0)
        with torch.no_grad():
            model.weight[0, 0] = 0.0
        micro_step()                                   # epoch 1 -> backup written
        assert (tmp_path / "s.zip").exists() and opt.local_epoch == 1
        good = model.weight.detach().clone()
        micro_step()                                   # epoch 2 (backup refreshed)
        good2, epoch2 = model.weight.detach().clone(), opt.local_epoch
        with torch.no_grad():
            model.weight[0, 3] = float("inf")
        cb.on_step_end(loss=1.0)                       # guard fires, state comes back from the file
        assert torch.isfinite(model.weight).all() and torch.equal(model.weight.detach(), good2)
        assert opt.local_epoch == epoch2 and not torch.equal(good, good2)
        opt.shutdown()
    finally:
        d0.shutdown()


def test_matchmaking_is_scheduled_ahead_of_the_epoch_and_overlaps_the_last_micro_batches():
    """--matchmaking_time (reference arguments.py:66-68): once the tracker expects the target batch within the
    matchmaking window, the group search runs in the background while micro-batches keep accumulating; the global
    step then finds its group waiting instead of opening a matchmaking window of its own."""
    d0 = _dht()
    d1 = _dht([str(m) for m in d0.get_visible_maddrs()])
    try:
        (m0, o0), (m1, o1) = _make(d0, "sched", bs=4, target=160), _make(d1, "sched", bs=4, target=160)
        waits, errors = {0: [], 1: []}, []

        def train(k, model, opt):
            try:
                while opt.local_epoch < 3:
                    for p in model.parameters():
                        p.grad = torch.ones_like(p) if p.grad is None else p.grad.add_(1.0)
                    before, t0 = opt.local_epoch, time.perf_counter()
                    opt.step()
                    if opt.local_epoch != before:
                        waits[k].append(time.perf_counter() - t0)
                    time.sleep(0.1)       # 20 micro-steps of 0.1 s per peer and epoch: the ETA is seconds away
            except Exception as e:  # pragma: no cover
                errors.append(e)

        ts = [threading.Thread(target=train, args=(0, m0, o0)), threading.Thread(target=train, args=(1, m1, o1))]
        [t.start() for t in ts]
        [t.join(120) for t in ts]
        assert not errors, errors
        assert o0.local_epoch >= 3 and o1.local_epoch >= 3
        # the first epoch has no speed estimate yet; afterwards rounds are pre-scheduled and were actually used
        assert o0.scheduled_rounds_used >= 1 and o1.scheduled_rounds_used >= 1, (o0.scheduled_rounds_used, o1.scheduled_rounds_used)
        # ... and a global step that found its group waiting is much shorter than the 1.5 s matchmaking window
        assert min(waits[0][1:] + waits[1][1:]) < 1.0, waits
        o0._finish_pending(); o1._finish_pending()
        assert torch.allclose(m0.weight, m1.weight, atol=1e-4)
        o0.shutdown(); o1.shutdown()
    finally:
        d0.shutdown(); d1.shutdown()
