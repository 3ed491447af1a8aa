#!/bin/bash
# Recipe for oracle/_ref: the UNMODIFIED reference package (kmerdb v0.9.6, pure Python) installed where it travels to
# the GPU box with the snapshot (oracle/_ref is git-ignored, not gpurun-ignored).  Run in the build container only
# (__graft_entry__.build() calls it when /root/reference is present).  Nothing of it enters the git history, and the
# product never imports it: only bench.py's CPU legs do (cpu_baseline / --impl reference), through tests/_bio_shim.
set -e
HERE=$(cd "$(dirname "$0")" && pwd)
REF=${1:-/root/reference}
[ -d "$REF/kmerdb" ] || { echo "no reference tree at $REF"; exit 0; }
rm -rf "$HERE/_ref"; mkdir -p "$HERE/_ref"
# the reference's own installer first (offline, no build isolation, no dependency resolution) ...
TMP=$(mktemp -d)
cp -r "$REF" "$TMP/src" 2>/dev/null && chmod -R u+w "$TMP/src"
if python -m pip install --quiet --no-index --no-build-isolation --no-deps --target "$HERE/_ref" "$TMP/src" > "$HERE/_ref/install.log" 2>&1 \
   && [ -f "$HERE/_ref/kmerdb/parse.py" ]; then
  echo "pip install --target oracle/_ref: ok" | tee -a "$HERE/_ref/install.log"
else
  # ... else the package directory as it lies in the reference tree (Python files only)
  echo "pip install failed or incomplete (see install.log): copying the package directory" | tee -a "$HERE/_ref/install.log"
  mkdir -p "$HERE/_ref/kmerdb"
  (cd "$REF/kmerdb" && find . -name '*.py' -o -name '*.yaml' -o -name '*.json' | cpio -pdm "$HERE/_ref/kmerdb" 2>/dev/null) || cp -r "$REF/kmerdb/." "$HERE/_ref/kmerdb/"
fi
rm -rf "$TMP"
( cd "$REF" && git rev-parse HEAD 2>/dev/null || true ) > "$HERE/_ref/REVISION"
ls "$HERE/_ref/kmerdb/parse.py" "$HERE/_ref/kmerdb/kmer.py"
