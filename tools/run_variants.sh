for v in default vA vB vC vD; do
  if [ $v = default ]; then unset NWB_LIB_PATH; else export NWB_LIB_PATH=tools/variants/$v.so; fi
  echo "== $v"; python tools/perf_field.py 10 2>&1 | grep "rays=1024"
done
