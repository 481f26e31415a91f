"""Raw-data preprocessing CLI (reference preprocess_data.py:1-47): FineWeb parquet ``text`` column -> documents of at
most 2000 characters, shuffled, split train/validation, dumped as ``{"train": [...], "validation": [...]}`` JSON."""
import json
import os
import random
from argparse import ArgumentParser


def get_args(argv=None):
    parser = ArgumentParser()
    parser.add_argument('data_path', type=str)
    parser.add_argument('output_path', type=str)
    parser.add_argument('--validation_parition', type=float, default=0.01)   # (sic) reference flag name
    parser.add_argument('--max_num_char', type=int, default=2000)
    parser.add_argument('--seed', type=int, default=None, help="seed the shuffle (the reference shuffles unseeded)")
    return parser.parse_args(argv)


def main(argv=None):
    import pandas as pd
    args = get_args(argv)
    assert os.path.exists(args.data_path), f"{args.data_path} not found"
    df = pd.read_parquet(args.data_path, columns=['text'])
    df['num_char'] = df['text'].apply(len)
    df = df[df['num_char'] <= args.max_num_char]
    extracted = df['text'].tolist()
    rng = random.Random(args.seed) if args.seed is not None else random
    rng.shuffle(extracted)
    train_num = int(len(extracted) * (1 - args.validation_parition))
    os.makedirs(os.path.dirname(args.output_path) or './', exist_ok=True)
    with open(args.output_path, 'w', encoding='utf-8') as f:
        json.dump({'train': extracted[:train_num], 'validation': extracted[train_num:]}, f, indent=2,
                  ensure_ascii=False)
    print(f"Training samples: {train_num}. Validation samples: {len(extracted) - train_num}. "
          f"Training chars: {sum(len(d) for d in extracted[:train_num])}. "
          f"Validation chars: {sum(len(d) for d in extracted[train_num:])}")


if __name__ == '__main__':
    main()
