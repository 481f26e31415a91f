"""Tokenizer trainer CLI (reference train_tokenizer.py:1-67): byte-level BPE with <BOS>/<EOS>/<UNK> = ids 0/1/2,
ByteLevel pre-tokenizer + decoder, saved as tokenizer.json, followed by a round-trip self-check."""
import json
import os
from argparse import ArgumentParser

from constants import BOS_TOKEN, EOS_TOKEN, UNK_TOKEN


def get_args(argv=None):
    parser = ArgumentParser()
    parser.add_argument("--data_path", "-d", type=str, required=True)
    parser.add_argument("--vocab_size", "-v", type=int, default=30000)
    parser.add_argument("--output_path", "-o", type=str, required=True)
    return parser.parse_args(argv)


def get_json_iterator(data_path: str, split: str):
    with open(data_path, "r") as f:
        data = json.load(f)
    for d in data[split]:
        yield d


def main(argv=None):
    from tokenizers import Tokenizer
    from tokenizers.decoders import ByteLevel as ByteLevelDecoder
    from tokenizers.models import BPE
    from tokenizers.pre_tokenizers import ByteLevel as ByteLevelPreTokenizer
    from tokenizers.trainers import BpeTrainer

    args = get_args(argv)
    tokenizer = Tokenizer(BPE(unk_token=UNK_TOKEN))
    tokenizer.pre_tokenizer = ByteLevelPreTokenizer()
    tokenizer.decoder = ByteLevelDecoder()
    trainer = BpeTrainer(vocab_size=args.vocab_size, special_tokens=[BOS_TOKEN, EOS_TOKEN, UNK_TOKEN])
    tokenizer.train_from_iterator(get_json_iterator(args.data_path, "train"), trainer=trainer)
    for name in (BOS_TOKEN, EOS_TOKEN, UNK_TOKEN):
        print(f"{name[1:-1]} token ID: {tokenizer.token_to_id(name)}")
    os.makedirs(os.path.dirname(args.output_path) or "./", exist_ok=True)   # bare filenames work (defect D10)
    tokenizer.save(args.output_path)
    print(f"Tokenizer saved to {args.output_path}")

    # round-trip self check
    tok = Tokenizer.from_file(args.output_path)
    for t in ["good morning", "hello world", "this is a test", "this is another test"]:
        decoded = tok.decode(tok.encode(t).ids).strip()
        assert t == decoded, f"'{t}' != '{decoded}'"


if __name__ == "__main__":
    main()
