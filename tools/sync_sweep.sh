for v in ${VARIANTS:-default s2 s3 s4}; do
  if [ $v = default ]; then unset BCP_B200_LIB; else export BCP_B200_LIB=/root/repo/oracle/_ref/libvar_$v.so; fi
  python bench.py --no-block --no-cpu-baseline --steps 10 --warmup 3 2>/dev/null > /tmp/b_$v.json
  python -c "import json; d=json.loads(open('/tmp/b_$v.json').read().strip().splitlines()[-1]); print('$v', round(d['value']/1e6,2), round(d['ms_per_step'],3))"
done
