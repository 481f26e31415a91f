"""Pre-tokenization CLI: text JSON -> token-id JSON, the file ``dataset.ShakespeareDataset`` reads.

Same flags and output schema as the reference's pre_tokenize.py:1-55 (``{split: [[ids...], ...], "special_ids":
{"<BOS>", "<EOS>", "<UNK>"}, "vocab_size": V}``); documents are encoded with ``Tokenizer.encode_batch`` in chunks
(the Rust tokenizer parallelises a batch across cores) instead of one ``encode`` call per document.
"""
import argparse
import json
import os
import sys

from constants import BOS_TOKEN, EOS_TOKEN, UNK_TOKEN

CHUNK = 2048      # documents per encode_batch call


def get_args(argv=None) -> argparse.Namespace:
    ap = argparse.ArgumentParser(description=__doc__.splitlines()[0])
    ap.add_argument("--input_file", "-i", required=True, help="text JSON written by preprocess_data.py")
    ap.add_argument("--output_file", "-o", required=True, help="token-id JSON to write")
    ap.add_argument("--tokenizer_file", "-t", required=True, help="tokenizer.json from train_tokenizer.py")
    ap.add_argument("--splits", "-s", nargs="+", default=["train", "validation"])
    return ap.parse_args(argv)


def encode_split(tokenizer, texts, label):
    """All documents of one split -> list of id lists, with a one-line progress report per chunk on stderr."""
    out = []
    for lo in range(0, len(texts), CHUNK):
        out.extend(enc.ids for enc in tokenizer.encode_batch(texts[lo:lo + CHUNK]))
        print(f"\r[{label}] {len(out)}/{len(texts)} documents", end="", file=sys.stderr, flush=True)
    if texts:
        print(file=sys.stderr)
    return out


def main(argv=None):
    from tokenizers import Tokenizer
    args = get_args(argv)
    for path in (args.input_file, args.tokenizer_file):
        if not os.path.isfile(path):
            raise SystemExit(f"pre_tokenize: {path} does not exist")
    with open(args.input_file, "r") as fh:
        corpus = json.load(fh)
    missing = [s for s in args.splits if s not in corpus]
    if missing:
        raise SystemExit(f"pre_tokenize: split(s) {missing} not in {args.input_file} (has {sorted(corpus)})")
    tok = Tokenizer.from_file(args.tokenizer_file)
    result = {}
    for split in args.splits:
        ids = encode_split(tok, corpus[split], split)
        n = len(ids)
        longest = max(map(len, ids), default=0)
        mean = sum(map(len, ids)) / n if n else 0.0
        print(f"{split}: {n} documents, longest {longest} tokens, mean {mean:.2f} tokens")
        result[split] = ids
    result["special_ids"] = {name: tok.token_to_id(name) for name in (BOS_TOKEN, EOS_TOKEN, UNK_TOKEN)}
    result["vocab_size"] = tok.get_vocab_size()
    os.makedirs(os.path.dirname(os.path.abspath(args.output_file)), exist_ok=True)
    with open(args.output_file, "w") as fh:
        json.dump(result, fh, ensure_ascii=False, separators=(",", ":"))


if __name__ == "__main__":
    main()
