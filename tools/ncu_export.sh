#!/bin/bash
# usage: tools/ncu_export.sh gpurun_out/<name>   (expects <name>.ncu-rep; writes text summaries, removes the rep)
# The .ncu-rep files are 15-30 MB each and gpurun_out/ is capped at 64 MiB: keep the text pages only.
R="$1"
[ -f "$R.ncu-rep" ] || exit 0
ncu -i "$R.ncu-rep" --page details > "$R.details.txt" 2>&1
ncu -i "$R.ncu-rep" --page raw --csv > "$R.raw.csv" 2>&1
python tools/ncu_source.py "$R.ncu-rep" 400 > "$R.source_segments.txt" 2>&1
ncu -i "$R.ncu-rep" --page source --csv 2>/dev/null | gzip -9 > "$R.source.csv.gz"
rm -f "$R.ncu-rep"
