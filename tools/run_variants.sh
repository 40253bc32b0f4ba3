# usage: bash tools/run_variants.sh "" name1 name2 ...   ("" = the product library)
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
for v in "$@"; do
  if [ -n "$v" ]; then export B200SPLAT_LIB=$PWD/3dgs_study_b200/variants/libb200splat_$v.so; fi
  echo "variant=$v"; python bench.py --steps 50 --warmup 5 --stages --no-cpu-baseline 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['stages_ms'])"
done
