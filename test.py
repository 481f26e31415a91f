"""Evaluation + greedy decoding entrypoint (reference test.py:23-172): validation loss of every checkpoint of this
rank (in iteration order) and greedy continuation of a few prompts at any TP size.

Fixes over the reference: the newest checkpoint is actually loaded before decoding (defect D2: it indexed the last
*character* of a path), prompts are not corrupted by the embedding at TP > 1 (D3), the mean is over batches (D5).
"""
import argparse
import os
from dataclasses import asdict

import torch
import torch.distributed as dist
import torch.multiprocessing as mp
import torch.nn.functional as F
import tqdm

import process_manager as pm
from constants import BOS_TOKEN, EOS_TOKEN, IGNORE_INDEX, ModelArgumments
from dataset import get_dataloader
from distributed_pytorch_from_scratch_b200.models.generation import greedy_generate
from distributed_pytorch_from_scratch_b200.utils.checkpoint import CKPT_RE, list_checkpoints, load_checkpoint
from distributed_pytorch_from_scratch_b200.utils.logging import make_summary_writer
from models.model import Transformer
from utils import init_dist_env, pick_backend, set_seed

PROMPTS = [
    "Nice to meet you, it's",
    "Great empire never falls, it only",
    "Your majesty, it's my duty ",
    "I shall be glad ",
    "What a glory to ",
    "Shame for the weak, it's",
    "The brave man ne",
    "Poor old man, it's",
]


def get_test_args(argv=None):
    parser = argparse.ArgumentParser()
    d = ModelArgumments()
    group = parser.add_argument_group("distributed")
    group.add_argument('--master_addr', type=str, default='localhost')
    group.add_argument('--master_port', type=str, default='23333')
    group.add_argument('--tp_size', type=int, default=2)
    group.add_argument('--backend', type=str, default=None, choices=[None, 'nccl', 'gloo'])
    group.add_argument('--comm', type=str, default='auto', choices=['auto', 'fused', 'nccl'])

    group = parser.add_argument_group("data")
    group.add_argument("--data_path", "-d", type=str, required=True)
    group.add_argument("--tokenizer_path", "-t", type=str, required=True)

    group = parser.add_argument_group("model")
    group.add_argument("--use_vallina_impl", action='store_true')
    parser.add_argument("--ckpt_dir", type=str, required=True)
    group.add_argument("--attn_dim", type=int, default=d.attn_dim)
    group.add_argument("--ffn_dim", type=int, default=d.ffn_dim)
    group.add_argument("--num_heads", type=int, default=d.num_heads)
    group.add_argument("--num_layers", type=int, default=d.num_layers)
    group.add_argument("--vocab_size", type=int, default=d.vocab_size)
    group.add_argument("--maxlen", type=int, default=d.maxlen)
    group.add_argument("--rope_theta", type=float, default=d.rope_theta)
    group.add_argument("--fp32", action='store_true', help="evaluate in fp32 instead of bf16")

    group = parser.add_argument_group("decode")
    group.add_argument("--max_decode_len", type=int, default=128)
    group.add_argument("--skip_decode", action='store_true')
    group.add_argument("--no_kv_cache", action='store_true',
                       help="decode like the reference: re-run the whole prefix for every new token "
                            "(default: incremental decoding with a per-layer K/V cache, same tokens)")

    group = parser.add_argument_group("other")
    group.add_argument("--random_seed", type=int, default=0)
    return parser.parse_args(argv)


def load_ckpt(model, ckpt_path: str, dtype: torch.dtype):
    model.eval()
    load_checkpoint(model, ckpt_path, dtype=dtype)


@torch.no_grad()
def calc_loss(model, ckpt_path: str, dataloader, dtype: torch.dtype, device, param_dtype=None) -> float:
    load_ckpt(model, ckpt_path, param_dtype if param_dtype is not None else dtype)
    pbar = tqdm.tqdm(range(len(dataloader)), desc=f"[TP Rank {pm.pgm.tp_rank}]: calc validation loss",
                     position=pm.pgm.tp_rank, disable=bool(os.environ.get("DPFS_NO_TQDM")))
    accum_loss, n_batches = 0., 0
    for batch in dataloader:
        ids, pos, tgt = (batch[k].to(device) for k in ('input_ids', 'position_ids', 'target_ids'))
        with torch.autocast(device.type, enabled=(dtype == torch.bfloat16), dtype=torch.bfloat16):
            logits = model(ids, pos)
        loss = F.cross_entropy(logits.float().view(-1, logits.size(-1)), tgt.view(-1), ignore_index=IGNORE_INDEX,
                               reduction='mean')
        accum_loss += loss.item()
        n_batches += 1
        pbar.update(1)
        pbar.set_postfix({'avg_loss': accum_loss / n_batches})
    pbar.close()
    return accum_loss / max(n_batches, 1)


@torch.no_grad()
def greedy_decode(model, tokens: torch.Tensor, eos_id: int, max_decode_len: int, dtype, device):
    """Full re-forward per generated token (no KV cache), like the reference (test.py:143-150)."""
    while True:
        position_ids = torch.arange(tokens.size(-1), dtype=torch.long, device=device).unsqueeze(0)
        with torch.autocast(device.type, enabled=(dtype == torch.bfloat16), dtype=torch.bfloat16):
            logits = model(tokens.clone(), position_ids)[0, -1]
        pred_token = int(logits.float().argmax(dim=-1).item())
        tokens = F.pad(tokens, (0, 1), mode="constant", value=pred_token)
        if pred_token == eos_id or tokens.size(-1) > max_decode_len:
            hit_eos = pred_token == eos_id
            if hit_eos:
                tokens = tokens[:, :-1]
            return tokens, hit_eos


def test(rank, args):
    set_seed(args.random_seed)
    init_dist_env(args, rank)
    backend = pick_backend(args)
    device = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")
    os.environ['DEVICE'] = device.type

    model_args = ModelArgumments(attn_dim=args.attn_dim, ffn_dim=args.ffn_dim, num_heads=args.num_heads,
                                 rope_theta=args.rope_theta, num_layers=args.num_layers, vocab_size=args.vocab_size,
                                 maxlen=args.maxlen)
    model = Transformer(**asdict(model_args))
    model.to(device)
    model.reset_parameters()
    model.eval()

    ckpt_paths = list_checkpoints(args.ckpt_dir, pm.pgm.tp_rank)
    if len(ckpt_paths) == 0:
        raise ValueError(f"[TP Rank {pm.pgm.tp_rank}]: No checkpoints found in {args.ckpt_dir}")
    print(f"[TP Rank {pm.pgm.tp_rank}]: Found {len(ckpt_paths)} checkpoints.")

    dtype = torch.float32 if args.fp32 else torch.bfloat16      # half precision for inference, like the reference
    os.environ["DTYPE"] = "float32" if args.fp32 else "bfloat16"
    model.compute_dtype = dtype
    # CPU: cast the parameters like the reference (test.py:101-103).  CUDA: keep fp32 masters; the fused path computes
    # in bf16 from the shadow weights, so the checkpoint is loaded without a lossy double cast.
    param_dtype = dtype if device.type == "cpu" else torch.float32
    model.to(param_dtype)

    dataloader = get_dataloader(args.data_path, 1, IGNORE_INDEX, split='validation', maxlen=model_args.maxlen,
                                shuffle=False)
    save_path = os.path.join(args.ckpt_dir, "val", f'tprank-{pm.pgm.tp_rank}_val.txt')
    os.makedirs(os.path.dirname(save_path), exist_ok=True)
    summary_writer = make_summary_writer(os.path.join(args.ckpt_dir, f"tprank-{rank}"))
    with open(save_path, 'a') as f:
        f.write("Ckpt -> Validation loss\n")
        for ckpt_path in ckpt_paths:
            iter_idx = int(CKPT_RE.search(os.path.basename(ckpt_path)).group(2))
            print(f"[TP Rank {pm.pgm.tp_rank}]: Evaluating iteration {iter_idx}")
            avg_loss = calc_loss(model, ckpt_path, dataloader, dtype, device, param_dtype)
            f.write(f"{ckpt_path} -> {avg_loss:.4f}\n")
            summary_writer.add_scalar("val/loss", avg_loss, iter_idx)
    summary_writer.close()

    if not args.skip_decode:
        from tokenizers import Tokenizer
        load_ckpt(model, ckpt_paths[-1], param_dtype)        # newest checkpoint
        tokenizer = Tokenizer.from_file(args.tokenizer_path)
        bos_id, eos_id = dataloader.dataset.bos, dataloader.dataset.eos
        assert tokenizer.token_to_id(BOS_TOKEN) == bos_id and tokenizer.token_to_id(EOS_TOKEN) == eos_id
        decoded = []
        for t in PROMPTS:
            t = t.strip()
            tokens = torch.tensor(tokenizer.encode(t).ids, dtype=torch.long, device=device).view(1, -1)
            tokens = F.pad(tokens, (1, 0), mode="constant", value=bos_id)
            if args.no_kv_cache:
                tokens, hit_eos = greedy_decode(model, tokens, eos_id, args.max_decode_len, dtype, device)
            else:
                tokens, hit_eos = greedy_generate(model, tokens, eos_id, args.max_decode_len, dtype=dtype)
            if not hit_eos:
                print(f"[TP Rank {pm.pgm.tp_rank}]: Maximum length reached. Stop decoding.")
            trans = tokenizer.decode(tokens[0, 1:].tolist()).strip()      # drop BOS
            assert t in trans, f"Prediction {trans} does not contain the input text {t}"
            decoded.append((t, trans[len(t):]))
        with open(save_path, 'a') as fp:
            print("\n\nInput texts -> Decoded texts", file=fp)
            for input_text, decoded_text in decoded:
                if rank == 0:
                    print(f"{input_text} -> {decoded_text}")
                print(f"{input_text} -> {decoded_text}", file=fp)

    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()
    pm.destroy_pgm()


def main(argv=None):
    args = get_test_args(argv)
    if "RANK" in os.environ and "WORLD_SIZE" in os.environ:
        args.master_addr = os.environ.get("MASTER_ADDR", args.master_addr)
        args.master_port = os.environ.get("MASTER_PORT", args.master_port)
        test(int(os.environ["RANK"]), args)
    else:
        mp.spawn(test, args=(args,), nprocs=args.tp_size, join=True)


if __name__ == '__main__':
    main()
