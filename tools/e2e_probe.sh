P='import json,sys; j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(sys.argv[1], "value", round(j["value"],1), "e2e", round(j["e2e"]["value"],1), j["e2e"]["link"])'
python bench.py --skip-sort --skip-micro 2>/dev/null | python -c "$P" quick1
python bench.py --skip-micro --skip-cpu-sort 2>/dev/null | python -c "$P" withsorts
python -m pytest tests/test_gpu_primitives.py -m gpu -x -q 2>&1 | tail -n 1
python bench.py --skip-sort --skip-micro 2>/dev/null | python -c "$P" after_pytest
python -c 'import __graft_entry__ as g; g.smoke()' 2>&1 | tail -n 1
python bench.py --skip-sort --skip-micro 2>/dev/null | python -c "$P" after_smoke
