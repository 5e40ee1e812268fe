"""Host-side data-parallel logic on CPU: world_size-2 gloo run of the bucketed gradient reduction
(GradBuckets), the backward-order parameter layout and the batch sharding.  The CUDA kernels are
not involved (no compute entry point is called without a GPU)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp
import torch.nn as nn


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from medvae_disentangled_multimodal_b200.train import GradBuckets, shard_batch
        torch.manual_seed(0)
        params = [nn.Parameter(torch.zeros(s)) for s in [(7, 3), (5,), (64, 9), (3,), (11, 11), (2,)]]
        gb = GradBuckets(params, bucket_bytes=256, group=None)   # several small buckets
        assert len(gb.bounds) >= 3
        assert all(p.grad.data_ptr() == gb.views[id(p)].data_ptr() for p in params)
        # two micro-batches: the first accumulates only, the second triggers the all-reduce
        g = torch.Generator().manual_seed(100 + rank)
        grads = [[torch.randn(p.shape, generator=g) for p in params] for _ in range(2)]
        gb.zero()
        for mb in range(2):
            gb.reduce_enabled = mb == 1
            gb.begin_round()
            for p, gr in zip(params, grads[mb]):
                gb.accumulate(p, gr)
            gb.flush()          # end of a backward pass: the batched small adds are applied and reported ready
        launched = gb.launched
        gb.finish()
        # expected: mean over ranks of the per-rank sums
        exp = []
        for i, p in enumerate(params):
            tot = torch.zeros(p.shape)
            for r in range(world):
                gg = torch.Generator().manual_seed(100 + r)
                both = [[torch.randn(q.shape, generator=gg) for q in params] for _ in range(2)]
                tot += both[0][i] + both[1][i]
            exp.append(tot / world)
        err = max(float((p.grad - e).abs().max()) for p, e in zip(params, exp))
        # a parameter that never reports ready must not dead-lock the other rank
        gb.zero()
        gb.reduce_enabled = True
        gb.begin_round()
        for p in params[:-1]:
            gb.accumulate(p, torch.ones(p.shape) * (rank + 1))
        gb.finish()
        err2 = max(float((p.grad - 1.5).abs().max()) for p in params[:-1])
        sl = shard_batch(8, rank, world)
        q.put((rank, err, err2, launched, len(gb.bounds), (sl.start, sl.stop)))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(120)
def test_grad_buckets_gloo_world2():
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=100) for _ in range(world)]
    for p in procs:
        p.join(timeout=30)
        assert p.exitcode == 0
    for rank, err, err2, launched, nb, sl in res:
        assert err < 1e-6 and err2 < 1e-6
        assert launched == nb            # every bucket was launched from accumulate(), none left to finish()
        assert sl == (rank * 4, rank * 4 + 4)


def test_backward_order_covers_every_parameter_once():
    import medvae_disentangled_multimodal_b200 as M
    from medvae_disentangled_multimodal_b200.train import backward_order
    m = M.ConditionalVAE(input_channels=3, latent_dim=4, hidden_channels=8, num_res_blocks=1, ch_mult=(1, 2),
                         attn_resolutions=[8], resolution=16)
    order = backward_order(m)
    assert len(order) == len(list(m.parameters())) == len({id(p) for p in order})
    names = {id(p): n for n, p in m.named_parameters()}
    seq = [names[id(p)] for p in order]
    assert seq[0] == "decoder.conv_out.weight"                      # last layer executed: first gradient ready
    assert seq.index("decoder.conv_in.weight") < seq.index("encoder.conv_out.weight")
    assert seq[-2:] == ["condition_proj.0.weight", "condition_proj.0.bias"]


def test_shard_batch_rejects_ragged():
    from medvae_disentangled_multimodal_b200.train import shard_batch
    with pytest.raises(ValueError):
        shard_batch(10, 0, 4)


def test_train_step_addresses_the_noise_by_global_sample_index(monkeypatch):
    """SURVEY 8e, host side: every micro-batch of a step uses ONE stream position, and model / encoder / decoder are told
    the position of the chunk's first sample in the global batch (rank * shard + chunk start); reset afterwards.  The
    device half (elem_base of mvae_reparam_kl_fwd) is checked in tests/test_zz_philox_gpu.py.  No GPU here: the
    optimizer and the model's forward are stand-ins, the chunk loop is the real TrainStep."""
    import medvae_disentangled_multimodal_b200 as M
    from medvae_disentangled_multimodal_b200 import optim, train

    class NoOptimizer:
        def __init__(self, params, **kw):
            self.params, self.step_state, self.lr, self.step_count = list(params), None, kw.get("lr"), 0

        def step(self, skip=None):
            self.step_count += 1

    monkeypatch.setattr(optim, "FusedAdamW", NoOptimizer)
    model = M.ConditionalVAE(input_channels=3, latent_dim=4, hidden_channels=8, ch_mult=(1, 2), num_res_blocks=1,
                             attn_resolutions=[], resolution=16, dropout=0.1)
    seen = []

    def forward(xs, cond=None):
        _seed, offset = model._next_rng()
        seen.append((offset, model._rng_sample_base, model.encoder._rng_sample_base, model.decoder._rng_sample_base,
                     xs.shape[0]))
        return {"reconstruction": torch.zeros(1, requires_grad=True), "posterior": None, "prior": None}

    model.forward = forward
    crit = lambda inputs, reconstructions, posteriors, priors: {"loss": reconstructions.sum() * 0.0}
    ts = train.TrainStep(model, crit, micro_batch=3)
    ts.buckets.rank = 2                                   # as rank 2 of a larger job, shard of 8 samples
    x, cond = torch.zeros(8, 3, 16, 16), torch.zeros(8, 12)
    ts.step(x, cond)
    ts.step(x, cond)
    ts.close()
    assert seen == [(1, 16, 16, 16, 3), (1, 19, 19, 19, 3), (1, 22, 22, 22, 2),
                    (2, 16, 16, 16, 3), (2, 19, 19, 19, 3), (2, 22, 22, 22, 2)]
    assert (model._rng_step, model._rng_sample_base, model.encoder._rng_sample_base) == (2, 0, 0)
    assert ts.optimizer.step_count == 2
