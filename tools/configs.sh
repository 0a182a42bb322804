# BASELINE.json configs C2..C5 shapes through bench.py (device-resident + e2e), one JSON line each
run() { echo "== $1"; shift; timeout 300 python bench.py --steps 5 --warmup 3 --windows-per-step 16 --no-cpu-baseline "$@" 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(json.dumps({k:d[k] for k in ('value','ms_per_step','gpu_launches')}), 'e2e', d['e2e']['value'], 'hits/step', d['e2e']['hits_per_step'], 'matcher', d['geometry']['matcher'], 'clk', d['clocks']['sm_mhz'])"; }
run C2 --format p2pkh --pattern '^1CatDog'
run C3 --format p2wpkh --pattern 'dead$'
run C4 --format ethereum --pattern '^0xdead'
run C5 --format p2pkh --pattern '^1D6vERHvZo8isjfDVdYXkoCtuMYxgaZEVG$'
run P2SH --format p2sh-p2wpkh --pattern '^3Cat'
run UNC --format p2pkh-uncompressed --pattern '^1Cat'
run DFA58 --format p2pkh --pattern 'Cat$'
