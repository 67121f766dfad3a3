#!/bin/bash
# sass_fn.sh OBJ PATTERN: SASS of the first function whose name matches PATTERN, one instruction per line
cuobjdump -sass "$1" | awk -v pat="$2" '/Function : /{f = ($0 ~ pat)} f' | grep -E "^\s+/\*[0-9a-f]{4}\*/" | sed 's/\/\* 0x[0-9a-f]* \*\///' | awk '{$1=$1};1'
