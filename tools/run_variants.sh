# usage: bash tools/run_variants.sh v1 v2 ...   (tools/variants/<v>.so; "default" = the in-tree library)
for v in "$@"; do
  if [ $v = default ]; then unset NWB_LIB_PATH; else export NWB_LIB_PATH=tools/variants/$v.so; fi
  echo "== $v"; python tools/perf_field.py 10 2>&1 | grep "rays=1024"
done
