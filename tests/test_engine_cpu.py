"""TrainStep / logging helpers on the CPU path (plain PyTorch optimizer, no CUDA graph): the public step API must work
without a GPU, advance the scheduler once per step and learn; the profiling helpers must degrade to no-ops."""
import torch
from torch.optim.lr_scheduler import OneCycleLR


def test_train_step_on_cpu_learns_and_steps_scheduler(tmp_path):
    import process_manager as pm
    from distributed_pytorch_from_scratch_b200.engine import TrainStep
    from distributed_pytorch_from_scratch_b200.utils.logging import (StepTimer, make_summary_writer, max_over_ranks,
                                                                     nvtx_range)
    from models.model import Transformer
    pm.init_pgm(1)
    torch.manual_seed(0)
    model = Transformer(attn_dim=64, ffn_dim=128, num_heads=4, num_layers=2, vocab_size=64, maxlen=32)
    model.reset_parameters()
    opt = torch.optim.Adam(model.parameters(), lr=3e-3)
    sched = OneCycleLR(opt, 3e-3, total_steps=20, pct_start=0.2)
    step = TrainStep(model, opt, sched, ignore_index=-1, use_graph=True)      # graph request is ignored on CPU
    assert step.use_graph is False and step.fused_opt is False
    ids = torch.randint(3, 64, (4, 16))
    pos = torch.arange(16).repeat(4, 1)
    tgt = torch.roll(ids, -1, dims=1)
    tgt[:, -1] = -1
    timer = StepTimer(torch.device("cpu"))
    timer.start()
    losses = []
    with nvtx_range("cpu-steps"):
        for _ in range(8):
            losses.append(float(step(ids, pos, tgt)))
    ms = timer.stop_ms()
    assert ms > 0 and max_over_ranks(ms, torch.device("cpu")) == ms
    assert step.steps == 8 and sched.last_epoch == 8
    assert losses[-1] < losses[0] - 0.3, losses
    # a second batch shape just works (static buffers are per shape)
    assert torch.isfinite(step(ids[:, :8], pos[:, :8], tgt[:, :8]))
    w = make_summary_writer(str(tmp_path / "tb"))
    w.add_scalar("train/ce_loss", losses[-1], 8)
    w.close()
    pm.destroy_pgm()


def test_shape_bucketing_changes_neither_loss_nor_update():
    """TrainStep(bucket=8) pads a 13-token batch to 16 positions (repeated last id, continued positions, ignored
    targets): with causal attention and an ignore-aware mean loss the step must be numerically the same step."""
    import process_manager as pm
    from distributed_pytorch_from_scratch_b200.engine import TrainStep
    from models.model import Transformer
    pm.init_pgm(1)
    ids = torch.randint(3, 64, (3, 13))
    pos = torch.arange(13).repeat(3, 1)
    tgt = torch.roll(ids, -1, dims=1)
    tgt[:, -1] = 1
    tgt[1, 9:] = -1                      # a shorter document inside the batch
    results = []
    for bucket in (0, 8):
        torch.manual_seed(0)
        model = Transformer(attn_dim=64, ffn_dim=128, num_heads=4, num_layers=2, vocab_size=64, maxlen=32)
        model.reset_parameters()
        opt = torch.optim.SGD(model.parameters(), lr=0.1)
        step = TrainStep(model, opt, None, ignore_index=-1, bucket=bucket)
        shape, _ = step.load_batch(ids, pos, tgt)
        assert shape == ((3, 16) if bucket else (3, 13))
        loss = float(step(ids, pos, tgt))
        results.append((loss, [p.detach().clone() for p in model.parameters()]))
    assert abs(results[0][0] - results[1][0]) < 1e-6
    for a, b in zip(results[0][1], results[1][1]):
        assert torch.allclose(a, b, atol=1e-6)
    # the cap: never past the RoPE table
    step = TrainStep(model, opt, None, ignore_index=-1, bucket=20)
    assert step._padded_len(25) == 32 and step._padded_len(13) == 20 and step._padded_len(40) == 40
    pm.destroy_pgm()
