"""train.py / test.py end to end on CPU + gloo at TP=2 with a synthetic corpus: checkpoints in the reference naming,
per-rank TensorBoard dirs, rotation, resume, per-checkpoint validation report and greedy decoding."""
import os
import subprocess
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MODEL = ["--attn_dim", "64", "--ffn_dim", "128", "--num_heads", "4", "--num_layers", "2", "--vocab_size", "1024",
         "--maxlen", "48"]


def _run(cmd, timeout=600):
    env = dict(os.environ, DPFS_NO_TQDM="1", PYTHONPATH=REPO)
    r = subprocess.run([sys.executable] + cmd, cwd=REPO, env=env, capture_output=True, text=True, timeout=timeout)
    assert r.returncode == 0, r.stdout[-3000:] + "\n" + r.stderr[-3000:]
    return r.stdout + r.stderr


def test_train_then_test_cli(tmp_path):
    from conftest import free_port
    save = str(tmp_path / "ck")
    out = _run(["train.py", "--tp_size", "2", "--synthetic_data", "--save_dir", save, "--max_steps", "6",
                "--warmup_steps", "2", "--log_interval", "2", "--save_interval", "2", "--reserv_last_n_ckpts", "2",
                "--batch_size", "2", "--backend", "gloo", "--master_port", str(free_port())] + MODEL)
    assert "Training finished (total steps: 6)" in out
    files = sorted(os.listdir(save))
    for r in (0, 1):
        ck = [f for f in files if f.startswith(f"tprank-{r}_iter-")]
        assert len(ck) == 2 and any("_iter-6_" in f for f in ck) and any("_iter-4_" in f for f in ck), files
        assert os.path.isdir(os.path.join(save, f"tprank-{r}"))            # per-rank TensorBoard dir
        assert os.path.exists(os.path.join(save, f"resume-tprank-{r}.pt"))
    # resume continues from step 6 to step 8
    out = _run(["train.py", "--tp_size", "2", "--synthetic_data", "--save_dir", save, "--max_steps", "8",
                "--warmup_steps", "2", "--log_interval", "2", "--save_interval", "2", "--batch_size", "2", "--resume",
                "--profile", "--backend", "gloo", "--master_port", str(free_port())] + MODEL)
    assert "Resumed from" in out and "at step 6" in out and "total steps: 8" in out
    # --profile reports throughput both ways: every padded position, and only positions that carry a target
    assert "[profile]" in out and "counting every padded position" in out and "is not padding" in out
    # a resume file that belongs to another step than every checkpoint on disk must be refused, not silently mixed
    import torch
    rp = os.path.join(save, "resume-tprank-0.pt")
    st = torch.load(rp, map_location="cpu", weights_only=False)
    st["step"] = 5
    torch.save(st, rp)
    env = dict(os.environ, DPFS_NO_TQDM="1", PYTHONPATH=REPO)
    bad = subprocess.run([sys.executable, "train.py", "--tp_size", "2", "--synthetic_data", "--save_dir", save,
                          "--max_steps", "9", "--warmup_steps", "2", "--batch_size", "2", "--resume", "--backend", "gloo",
                          "--master_port", str(free_port())] + MODEL, cwd=REPO, env=env, capture_output=True, text=True,
                         timeout=600)
    assert bad.returncode != 0 and "refusing to mix them" in (bad.stdout + bad.stderr)
    st["step"] = 8
    torch.save(st, rp)
    # evaluation + greedy decode at TP=2
    data = os.path.join(save, "synthetic_tokens.json")
    out = _run(["test.py", "--tp_size", "2", "--ckpt_dir", save, "-d", data, "-t", "tokenizer/tokenizer.json",
                "--max_decode_len", "12", "--fp32", "--backend", "gloo", "--master_port", str(free_port())] + MODEL)
    rep = open(os.path.join(save, "val", "tprank-0_val.txt")).read()
    assert "Ckpt -> Validation loss" in rep and "_iter-8_" in rep and "Input texts -> Decoded texts" in rep
    assert "Nice to meet you, it's ->" in rep
    rep1 = open(os.path.join(save, "val", "tprank-1_val.txt")).read()
    assert rep.split("Input texts")[1] == rep1.split("Input texts")[1]     # every rank decodes the same text
    # the default decode keeps a K/V cache; the reference-style full re-forward must produce the same text
    kv_text = rep.split("Input texts")[1]
    os.remove(os.path.join(save, "val", "tprank-0_val.txt"))
    _run(["test.py", "--tp_size", "2", "--ckpt_dir", save, "-d", data, "-t", "tokenizer/tokenizer.json",
          "--max_decode_len", "12", "--fp32", "--no_kv_cache", "--backend", "gloo", "--master_port",
          str(free_port())] + MODEL)
    rep = open(os.path.join(save, "val", "tprank-0_val.txt")).read()
    assert rep.split("Input texts")[1] == kv_text
