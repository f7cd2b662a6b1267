#!/bin/bash
# dump the SASS of one kernel of libugdet.so:  tools/sass.sh <mangled-name-regex> > out.sass
cuobjdump -sass "$(dirname "$0")/../ugproj-detector_b200/libugdet.so" 2>/dev/null | awk -v pat="$1" '/Function :/ {on = ($0 ~ pat)} on' | grep -E "^\s+/\*[0-9a-f]{4,}\*/" | sed 's/\/\* 0x[0-9a-f]* \*\///'
