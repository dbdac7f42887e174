#!/bin/sh
# Turns the device headers that NVRTC needs at run time (zcu_graph.cu) into C string literals.
# usage: embed_headers.sh name:file ... > build/zcu_embedded_headers.inc
for pair in "$@"; do
    name=${pair%%:*}
    file=${pair#*:}
    printf 'const char* const kEmbedded_%s =\n' "$name"
    sed -e 's/\\/\\\\/g' -e 's/"/\\"/g' -e 's/^/    "/' -e 's/$/\\n"/' "$file"
    printf '    ;\n'
done
