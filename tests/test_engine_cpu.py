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
