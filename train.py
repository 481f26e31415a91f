"""Training entrypoint — same CLI, launch model, logging and checkpoint behaviour as the reference (train.py:25-151)
on the B200-native stack.

    python train.py --tp_size N --data_path tokens.json [--bf16] ...        # self-spawns N ranks (mp.spawn)
    torchrun --nproc-per-node N train.py --tp_size N ...                    # also works (uses RANK/LOCAL_RANK)

Every reference flag keeps its name and default.  Added: model dims (the reference hard-codes
``ModelArgumments``), ``--backend {nccl,gloo}``, ``--synthetic_data``, ``--comm {auto,fused,nccl}``,
``--sequence_parallel``, ``--cuda_graph``, ``--profile``, ``--resume``.
"""
import math
import os
from argparse import ArgumentParser, Namespace
from dataclasses import asdict

import torch
import torch.distributed as dist
import torch.multiprocessing as mp
import torch.optim as optim
import tqdm

import process_manager as pm
from constants import IGNORE_INDEX, ModelArgumments
from dataset import get_dataloader, write_synthetic_corpus
from distributed_pytorch_from_scratch_b200.engine import TrainStep
from distributed_pytorch_from_scratch_b200.ops.optim import FusedAdam
from distributed_pytorch_from_scratch_b200.utils.checkpoint import (list_checkpoints, load_checkpoint,
                                                                    load_resume_state, save_checkpoint,
                                                                    save_resume_state)
from distributed_pytorch_from_scratch_b200.utils.logging import StepTimer, make_summary_writer, max_over_ranks
from models.model import Transformer
from utils import init_dist_env, pick_backend, set_seed


def get_train_args(argv=None):
    parser = ArgumentParser()
    d = ModelArgumments()

    group = parser.add_argument_group("distributed")
    group.add_argument('--tp_size', type=int, default=2)
    group.add_argument('--master_addr', type=str, default='localhost')
    group.add_argument('--master_port', type=str, default='25555')
    group.add_argument('--backend', type=str, default=None, choices=[None, 'nccl', 'gloo'],
                       help="default: nccl when CUDA is available, else gloo (CPU)")
    group.add_argument('--comm', type=str, default='auto', choices=['auto', 'fused', 'nccl'],
                       help="TP collectives: in-kernel NVLink (fused) or NCCL; auto = fused when available")
    group.add_argument('--sequence_parallel', action='store_true',
                       help="Megatron sequence parallelism: reduce-scatter / all-gather instead of all-reduce")

    group = parser.add_argument_group("training")
    group.add_argument("--lr", type=float, default=3e-4)
    group.add_argument("--warmup_steps", type=int, default=2000)
    group.add_argument("--max_steps", type=int, default=20000)
    group.add_argument("--log_interval", type=int, default=100)
    group.add_argument("--save_interval", type=int, default=1000)
    group.add_argument("--save_dir", type=str, default='./checkpoints')
    group.add_argument("--reserv_last_n_ckpts", type=int, default=-1)
    group.add_argument("--batch_size", "-b", type=int, default=32)
    group.add_argument("--bf16", action='store_true', help="Whether to use automatic mixed precision or not.")
    group.add_argument("--cuda_graph", action='store_true', help="replay the whole step as a CUDA graph")
    group.add_argument("--bucket_len", type=int, default=128,
                       help="with --cuda_graph: pad sequence lengths up to a multiple of this (0 = exact shapes)")
    group.add_argument("--resume", action='store_true',
                       help="continue from the newest checkpoint + resume-tprank-{r}.pt in --save_dir")
    group.add_argument("--profile", action='store_true', help="print device-timed step time / tokens per second")

    group = parser.add_argument_group("model (defaults = constants.ModelArgumments)")
    group.add_argument("--attn_dim", type=int, default=d.attn_dim)
    group.add_argument("--ffn_dim", type=int, default=d.ffn_dim)
    group.add_argument("--num_heads", type=int, default=d.num_heads)
    group.add_argument("--num_layers", type=int, default=d.num_layers)
    group.add_argument("--vocab_size", type=int, default=d.vocab_size)
    group.add_argument("--maxlen", type=int, default=d.maxlen)
    group.add_argument("--rope_theta", type=float, default=d.rope_theta)

    group = parser.add_argument_group("data")
    group.add_argument("--data_path", "-d", type=str, default=None)
    group.add_argument("--synthetic_data", action='store_true',
                       help="generate a synthetic token corpus with the reference JSON schema (no network needed)")

    group = parser.add_argument_group("other")
    group.add_argument('--random_seed', type=int, default=0)
    group.add_argument('--use_vallina_impl', action='store_true',
                       help="Train the single-device vanilla oracle instead of the TP model (requires tp_size 1).")

    args = parser.parse_args(argv)
    if args.data_path is None and not args.synthetic_data:
        parser.error("--data_path/-d is required unless --synthetic_data is given")
    return args


def _prepare_data(args: Namespace) -> str:
    if args.data_path is not None and os.path.exists(args.data_path):
        return args.data_path
    assert args.synthetic_data, f"data file {args.data_path} not found"
    path = args.data_path or os.path.join(args.save_dir, "synthetic_tokens.json")
    if not os.path.exists(path):
        write_synthetic_corpus(path, vocab_size=args.vocab_size, n_train=max(4 * args.batch_size, 64), n_val=8,
                               min_len=max(8, args.maxlen // 4), max_len=args.maxlen - 1, seed=args.random_seed)
    return path


def train(rank: int, args: Namespace):
    set_seed(args.random_seed)
    init_dist_env(args, rank)
    backend = pick_backend(args)
    device = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")
    os.environ['DEVICE'] = device.type
    if args.bf16:
        print("Enable bf16 training")
        os.environ['DTYPE'] = 'bfloat16'
    else:
        print("Disable bf16 training")
        os.environ['DTYPE'] = 'float32'
    dtype = torch.bfloat16 if args.bf16 else torch.float32

    model_args = ModelArgumments(attn_dim=args.attn_dim, ffn_dim=args.ffn_dim, num_heads=args.num_heads,
                                 rope_theta=args.rope_theta, num_layers=args.num_layers, vocab_size=args.vocab_size,
                                 maxlen=args.maxlen)
    if args.use_vallina_impl:
        assert args.tp_size == 1, "--use_vallina_impl trains the single-device oracle: use --tp_size 1"
        from models.model import VallinaTransformer
        model = VallinaTransformer(**asdict(model_args))
    else:
        model = Transformer(**asdict(model_args), sequence_parallel=args.sequence_parallel)
        model.compute_dtype = dtype
    model.to(device)
    model.reset_parameters()        # re-initialize parameters (necessary for the tensor-parallel Transformer)
    model.train()

    if pm.pgm.tp_rank == 0:
        print(model)
        num_param = sum(p.numel() for p in model.parameters() if p.requires_grad)
        print(f"Number of parameters: {num_param / 1e6:.4f} million")

    if rank == 0:
        data_path = _prepare_data(args)
    if dist.is_initialized():
        dist.barrier()
    data_path = _prepare_data(args)
    dataloader = get_dataloader(data_path, args.batch_size, IGNORE_INDEX, split='train', maxlen=model_args.maxlen,
                                shuffle=True)
    assert dataloader.dataset.vocab_size == model_args.vocab_size, "vocab size of dataset and model should be the same"

    if dist.is_initialized():
        dist.barrier()
    fused_model = isinstance(model, Transformer)
    if fused_model:
        optimizer = FusedAdam.for_model(model, lr=args.lr)
    else:
        optimizer = optim.Adam(model.parameters(), lr=args.lr)
    scheduler = optim.lr_scheduler.OneCycleLR(optimizer, args.lr, args.max_steps,
                                              pct_start=args.warmup_steps / args.max_steps)
    optimizer._opt_called = True
    summary_writer = make_summary_writer(os.path.join(args.save_dir, f"tprank-{rank}"))   # every rank logs

    start_step, accum_loss = 0, 0.
    if args.resume:
        # Weights come from the tprank-{r}_iter-{n}_... checkpoint, optimizer / scheduler / step from the resume file
        # written right after it.  They must belong to the SAME step (a crash between the two writes, or a shard
        # copied next to an old resume file, would pair step-N weights with step-M Adam moments): pick the checkpoint
        # whose iteration equals the resume file's step, refuse otherwise.  The data order restarts at the beginning
        # of an epoch on resume (the loader position is not part of the state).
        ckpts = list_checkpoints(args.save_dir, pm.pgm.tp_rank)
        if ckpts:
            import re as _re
            st = load_resume_state(args.save_dir, pm.pgm.tp_rank, optimizer, scheduler)
            chosen = ckpts[-1]
            if st is not None:
                by_iter = {int(_re.search(r"_iter-(\d+)_", os.path.basename(c)).group(1)): c for c in ckpts}
                if st["step"] not in by_iter:
                    raise RuntimeError(f"resume-tprank-{pm.pgm.tp_rank}.pt is from step {st['step']} but the checkpoints "
                                       f"in {args.save_dir} are from steps {sorted(by_iter)}: refusing to mix them")
                chosen = by_iter[st["step"]]
                start_step, accum_loss = st["step"], st["accum_loss"]
            load_checkpoint(model, chosen)
            print(f"[TP rank {rank}]: Resumed from {chosen} at step {start_step}")

    if fused_model:
        # with --cuda_graph, ragged batches are padded up to multiples of --bucket_len so that a handful of captured
        # graphs cover every batch (no loss / gradient change: padded positions are causal-masked and ignored)
        bucket = args.bucket_len if args.cuda_graph else 0
        n_shapes = (model_args.maxlen + bucket - 1) // bucket + 1 if bucket > 0 else 4
        step_fn = TrainStep(model, optimizer, scheduler, ignore_index=IGNORE_INDEX, use_graph=args.cuda_graph,
                            bucket=bucket, max_graphs=max(4, min(n_shapes, 32)))
    else:
        step_fn = None

    tag = f"TP-{pm.pgm.tp_rank}" if not args.use_vallina_impl else "vanilla"
    pbar = tqdm.tqdm(range(args.max_steps), desc=f"Training-[{tag}]", position=rank, initial=start_step,
                     disable=bool(os.environ.get("DPFS_NO_TQDM")))
    max_epoch = math.ceil(args.max_steps / max(len(dataloader), 1))
    accum_dev = torch.zeros((), dtype=torch.float64, device=device)   # running loss sum stays on the device
    n_step = start_step
    timer = StepTimer(device)
    tokens_since, real_since, steps_since = 0, 0, 0     # padded positions / positions with a real target / steps
    timer.start()

    if dist.is_initialized():
        dist.barrier()
    done = n_step >= args.max_steps
    for epoch in range(max_epoch):
        if done:
            break
        for batch in dataloader:
            if step_fn is not None:
                loss = step_fn(batch['input_ids'], batch['position_ids'], batch['target_ids'])
            else:
                ids, pos, tgt = (batch[k].to(device) for k in ('input_ids', 'position_ids', 'target_ids'))
                with torch.autocast(device.type, enabled=args.bf16, dtype=torch.bfloat16):
                    logits = model(ids, pos)
                    loss = torch.nn.functional.cross_entropy(logits.float().view(-1, logits.size(-1)), tgt.view(-1),
                                                             ignore_index=IGNORE_INDEX, reduction='mean')
                optimizer.zero_grad()
                loss.backward()
                optimizer.step()
                scheduler.step()
                loss = loss.detach()
            accum_dev += loss.double()
            n_step += 1
            tokens_since += batch['input_ids'].numel()
            if args.profile:     # counted on the host copy of the batch: no device sync
                real_since += int((batch['target_ids'] != IGNORE_INDEX).sum())
            steps_since += 1
            pbar.update(1)
            if n_step % args.log_interval == 0:
                avg_loss = accum_loss + accum_dev.item()       # one host sync per log interval, not per step
                avg_loss /= n_step
                lr = optimizer.param_groups[0]['lr']
                pbar.set_postfix({'avg_loss': avg_loss})
                print(f"[TP rank {rank}]: Step {n_step}/{args.max_steps} -> Avg Loss {avg_loss:.4f}, Lr {lr:.8f}")
                summary_writer.add_scalar('train/ce_loss', avg_loss, n_step)
                summary_writer.add_scalar('train/lr', lr, n_step)
                if device.type == "cuda":
                    reserved_memory = torch.cuda.memory_reserved(torch.cuda.current_device()) / 1024 ** 3   # GiB
                    summary_writer.add_scalar(f'used_gpu_memory/tprank-{rank}', reserved_memory, n_step)
                if args.profile:
                    ms = max_over_ranks(timer.stop_ms(), device) / max(steps_since, 1)
                    tps = tokens_since / max(steps_since, 1) / ms * 1e3
                    tps_real = real_since / max(steps_since, 1) / ms * 1e3
                    if rank == 0:
                        print(f"[profile] {ms:.2f} ms/step (device-timed, max over ranks), {tps:,.0f} tokens/s counting "
                              f"every padded position, {tps_real:,.0f} tokens/s counting only positions with a target "
                              f"({100.0 * real_since / max(tokens_since, 1):.1f}% of the batch is not padding)")
                    summary_writer.add_scalar('train/tokens_per_sec', tps, n_step)
                    summary_writer.add_scalar('train/nonpad_tokens_per_sec', tps_real, n_step)
                    tokens_since, real_since, steps_since = 0, 0, 0
                    timer.start()
            if n_step % args.save_interval == 0:
                avg_loss = (accum_loss + accum_dev.item()) / n_step
                save_checkpoint(model, args.save_dir, rank, n_step, avg_loss, keep_last_n=args.reserv_last_n_ckpts)
                save_resume_state(args.save_dir, rank, n_step, optimizer, scheduler, accum_loss + accum_dev.item())
                if dist.is_initialized():
                    dist.barrier()
            if n_step >= args.max_steps:
                done = True
                break
        if pm.pgm.tp_rank == 0:
            print(f"Epoch {epoch + 1}/{max_epoch} finished.")
        if done:
            print(f"[TP rank {rank}]: Training finished (total steps: {n_step}). Exiting...")
            break
        if device.type == "cuda":
            torch.cuda.empty_cache()

    pbar.close()
    summary_writer.close()
    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()
    pm.destroy_pgm()


def main(argv=None):
    args = get_train_args(argv)
    if "RANK" in os.environ and "WORLD_SIZE" in os.environ:      # launched by torchrun
        assert int(os.environ["WORLD_SIZE"]) == args.tp_size
        args.master_addr = os.environ.get("MASTER_ADDR", args.master_addr)
        args.master_port = os.environ.get("MASTER_PORT", args.master_port)
        train(int(os.environ["RANK"]), args)
    else:
        mp.spawn(train, args=(args,), nprocs=args.tp_size, join=True)


if __name__ == '__main__':
    main()
