#!/bin/bash
# Static opcode mix of one kernel in the built library (development aid).  usage: tools/dev/sass_mix.sh <regex>
cuobjdump -sass melissa_b200/libmelissa_b200.so 2>/dev/null | awk -v pat="$1" '
/Function :/ {on = ($0 ~ pat)} on && /^\s+\/\*[0-9a-f]+\*\// {print}' | sed -E 's/^\s+\/\*[0-9a-f]+\*\/\s+//; s/^@!?U?P[0-9T]+\s+//; s/ .*//; s/\..*//; s/;//' | sort | uniq -c | sort -rn | head -${2:-30}
