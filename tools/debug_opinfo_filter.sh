#!/bin/bash
# prints only the samples where b200 fails and the torch executor passes
python tools/debug_opinfo.py "$@" 2>&1 | grep -v Warning | awk '/ sample /{hdr=$0; next} /^   b200 :/{b=$0; next} /^   torch:/{ if (b !~ /: ok$/ && $0 ~ /: ok$/) {print hdr; print b; print $0; show=1} else show=0; next} { if (show) print }' | cut -c1-400
