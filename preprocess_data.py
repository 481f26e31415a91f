"""Raw-data preprocessing CLI: one parquet shard with a ``text`` column (FineWeb) -> ``{"train": [...],
"validation": [...]}`` JSON of short documents.

Same positional arguments, flags and output schema as the reference's preprocess_data.py:1-47 (including its
``--validation_parition`` spelling); ``--seed`` and ``--max_num_char`` are additions.  The text column is filtered with
an Arrow compute kernel and shuffled through an index permutation; no pandas frame is built.
"""
import argparse
import json
import os
import random


def get_args(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.splitlines()[0])
    ap.add_argument("data_path", help="input .parquet file")
    ap.add_argument("output_path", help="output .json file")
    ap.add_argument("--validation_parition", type=float, default=0.01,
                    help="fraction of documents held out for validation (flag name as in the reference)")
    ap.add_argument("--max_num_char", type=int, default=2000, help="drop documents longer than this")
    ap.add_argument("--seed", type=int, default=None, help="seed of the shuffle (unseeded by default, as the reference)")
    return ap.parse_args(argv)


def short_documents(parquet_path: str, max_chars: int):
    import pyarrow.compute as pc
    import pyarrow.parquet as pq
    col = pq.read_table(parquet_path, columns=["text"]).column("text")
    return col.filter(pc.less_equal(pc.utf8_length(col), max_chars)).to_pylist()


def main(argv=None):
    args = get_args(argv)
    if not os.path.isfile(args.data_path):
        raise SystemExit(f"preprocess_data: {args.data_path} does not exist")
    docs = short_documents(args.data_path, args.max_num_char)
    order = list(range(len(docs)))
    (random.Random(args.seed) if args.seed is not None else random).shuffle(order)
    n_val = len(docs) - int(len(docs) * (1.0 - args.validation_parition))
    splits = {"train": [docs[i] for i in order[n_val:]], "validation": [docs[i] for i in order[:n_val]]}
    os.makedirs(os.path.dirname(os.path.abspath(args.output_path)), exist_ok=True)
    with open(args.output_path, "w", encoding="utf-8") as fh:
        json.dump(splits, fh, ensure_ascii=False, indent=2)
    for name, items in splits.items():
        print(f"{name}: {len(items)} documents, {sum(map(len, items))} characters")


if __name__ == "__main__":
    main()
