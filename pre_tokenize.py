"""Pre-tokenization CLI (reference pre_tokenize.py:1-55): text JSON -> token-id JSON with ``special_ids`` and
``vocab_size`` appended — exactly the file dataset.ShakespeareDataset reads."""
import json
import os
from argparse import ArgumentParser, Namespace

import tqdm

from constants import BOS_TOKEN, EOS_TOKEN, UNK_TOKEN


def get_args(argv=None) -> Namespace:
    parser = ArgumentParser()
    parser.add_argument('--input_file', '-i', type=str, required=True)
    parser.add_argument('--output_file', '-o', type=str, required=True)
    parser.add_argument('--tokenizer_file', '-t', type=str, required=True)
    parser.add_argument('--splits', '-s', type=str, nargs='+', default=['train', 'validation'])
    return parser.parse_args(argv)


def main(argv=None):
    from tokenizers import Tokenizer
    args = get_args(argv)
    assert os.path.exists(args.input_file), f'Input file {args.input_file} not found'
    with open(args.input_file, 'r') as f:
        datas = json.load(f)
    assert all(s in datas for s in args.splits), f'Expected splits {args.splits}, but found {list(datas.keys())}'
    assert os.path.exists(args.tokenizer_file), f'Tokenizer file {args.tokenizer_file} not found'
    tokenizer = Tokenizer.from_file(args.tokenizer_file)
    token_data = {}
    for split in args.splits:
        token_data[split], lens = [], []
        for text in tqdm.tqdm(datas[split], desc=f'Tokenizing {split}'):
            ids = tokenizer.encode(text).ids
            token_data[split].append(ids)
            lens.append(len(ids))
        print(f"Split: {split} -> Number of samples: {len(lens)}. Max num_tokens: {max(lens) if lens else 0}. "
              f"Avg num_tokens: {(sum(lens) / len(lens)) if lens else 0:.2f}.")
    token_data["special_ids"] = {t: tokenizer.token_to_id(t) for t in (BOS_TOKEN, EOS_TOKEN, UNK_TOKEN)}
    token_data["vocab_size"] = tokenizer.get_vocab_size()
    os.makedirs(os.path.dirname(args.output_file) or "./", exist_ok=True)
    with open(args.output_file, 'w') as f:
        json.dump(token_data, f, ensure_ascii=False)


if __name__ == '__main__':
    main()
