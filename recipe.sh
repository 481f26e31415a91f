#!/bin/bash
# End-to-end recipe (reference recipe.sh:1-125): data -> tokenizer -> token ids -> train TP1/2/4/8 -> test.
# Every step is idempotent (skipped when its output exists).  There is no network on the build / GPU boxes, so when
# the FineWeb parquet is absent (and cannot be downloaded) the recipe falls back to a synthetic token corpus with the
# same JSON schema (SYNTHETIC=1 forces it).
set -e

ori_data_path="${ORI_DATA_PATH:-./fineweb/CC-MAIN-2023-50/000_00000.parquet}"
json_data_path="${JSON_DATA_PATH:-data.json}"
token_data_path="${json_data_path%.*}_tokens.json"
vocab=1024
tokenizer_path="./tokenizer/tokenizer.json"
root_ckpt_dir="${ROOT_CKPT_DIR:-./.checkpoints}"
tp_sizes="${TP_SIZES:-1 2 4 8}"
train_flags="${TRAIN_FLAGS:---bf16 --cuda_graph}"
step_idx=0
step() { step_idx=$((step_idx+1)); echo -e "Step ${step_idx}: $1\n"; }

if [ "${SYNTHETIC:-0}" != "1" ]; then
  step "Downloading raw data..."
  if [ ! -f "${ori_data_path}" ]; then
    data_url="https://huggingface.co/datasets/HuggingFaceFW/fineweb/resolve/main/data/CC-MAIN-2023-50/000_00000.parquet?download=true"
    mkdir -p "$(dirname "${ori_data_path}")"
    curl -L --fail --max-time 600 "${data_url}" -o "${ori_data_path}" || { echo "download failed -> synthetic corpus"; SYNTHETIC=1; rm -f "${ori_data_path}"; }
  else
    echo -e "raw data path exists, skip downloading\n"
  fi
fi

if [ "${SYNTHETIC:-0}" == "1" ]; then
  step "Generating a synthetic token corpus (reference JSON schema)..."
  if [ ! -f "${token_data_path}" ]; then
    python3 -c "from dataset import write_synthetic_corpus; write_synthetic_corpus('${token_data_path}', vocab_size=${vocab}, n_train=2048, n_val=64)"
  fi
else
  step "Preprocessing data..."
  [ -f "${json_data_path}" ] || python3 ./preprocess_data.py "${ori_data_path}" "${json_data_path}"
  step "Train tokenizers..."
  [ -f "${tokenizer_path}" ] || python3 ./train_tokenizer.py --data_path "${json_data_path}" --vocab_size "${vocab}" --output_path "${tokenizer_path}"
  step "Tokenizing data..."
  [ -f "${token_data_path}" ] || python3 ./pre_tokenize.py -i "${json_data_path}" -o "${token_data_path}" -t "${tokenizer_path}"
fi

port=23330
for tp in ${tp_sizes}; do
  step "Train TP-${tp} model..."
  ckpt_dir="${root_ckpt_dir}/tp${tp}"
  if [ ! -d "${ckpt_dir}" ]; then
    python3 ./train.py --tp_size "${tp}" --save_dir "${ckpt_dir}" --data_path "${token_data_path}" \
      --master_addr 127.0.0.1 --master_port "${port}" ${train_flags} ${EXTRA_TRAIN_FLAGS}
  else
    echo -e "TP-${tp} model ckpt exists, skip training\n"
  fi
  port=$((port+1))
done

port=23340
for tp in ${tp_sizes}; do
  step "Test TP-${tp} model..."
  ckpt_dir="${root_ckpt_dir}/tp${tp}"
  if [ ! -d "${ckpt_dir}/val" ]; then
    python3 ./test.py --tp_size "${tp}" --ckpt_dir "${ckpt_dir}" --data_path "${token_data_path}" \
      --tokenizer_path "${tokenizer_path}" --master_addr 127.0.0.1 --master_port "${port}" ${EXTRA_TEST_FLAGS}
  else
    echo -e "TP-${tp} model test result exists, skip testing\n"
  fi
  port=$((port+1))
done
