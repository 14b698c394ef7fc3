for e in neohookean_c neohookean_i arap; do
  python bench.py --steps 5 --warmup 2 --energy $e --nx 201 --ny 51 --nz 41 --no-solves --no-cpu --no-energies 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); k=d['roofline']['kernel_ms']
print('$e', 'value %.4g'%d['value'], 'jac %.3f'%k['k_fused_jacobian']['ms_per_sweep'], 'eval %.3f'%k['k_fused_eval']['ms_per_sweep'], 'step %.3f'%k['k_fused_step']['ms_per_sweep'])"
done
