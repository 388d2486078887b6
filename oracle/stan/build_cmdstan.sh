#!/bin/bash
# TEST INFRASTRUCTURE: build the reference's vendored CmdStan 2.20.0 in a writable copy (/root/reference is
# read-only).  ~90 s.  Used only by oracle/stan/make_golden.py to GENERATE tests/golden/*_stan.npz.
set -e
REF=${REF:-/root/reference/stochaskell}
DST=${CMDSTAN:-/tmp/cmdstan}
if [ ! -x "$DST/bin/stanc" ]; then
  rm -rf "$DST"; cp -r "$REF/cmdstan" "$DST"; chmod -R u+w "$DST"
  make -C "$DST" build -j"$(nproc)" > "$DST/build.log" 2>&1
fi
echo "cmdstan ready at $DST"
