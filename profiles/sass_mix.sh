#!/bin/bash
# usage: sass_mix.sh <object-or-so> <function-substring>   -- instruction count and opcode mix of one kernel
cuobjdump -sass "$1" 2>/dev/null | awk -v pat="$2" '
/Function : /{f = (index($0, pat) > 0)}
f && /^ +\/\*[0-9a-f]+\*\/ +[A-Z@!]/ { op=$2; if (op ~ /^@/) op=$3; sub(/\..*/,"",op); sub(/;/,"",op); c[op]++; n++ }
END{ print n, "instructions"; for (o in c) print c[o], o }' | sort -rn | head -${3:-30}
