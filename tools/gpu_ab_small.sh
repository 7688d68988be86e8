for name in base sb16 b32; do
  if [ $name = base ]; then unset DRT_LIB_PATH; else export DRT_LIB_PATH=$PWD/dustraytracer_b200/ab/libdrt_$name.so; fi
  echo "# lib $name"
  timeout 600 python tools/sweep.py --scenes room,sunshadow_test,mc_transparency_test,bvhsplit_test --spp 64 --variants -1 --blocks 128 --reps 3 2>/dev/null | grep "^{" | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print('  %-22s nodes %6d %9.3f ms %9.1f Mrays/s'%(d['scene'],d['wide_nodes'],d['ms'],d['Mrays_s']))"
done
