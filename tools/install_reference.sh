#!/bin/bash
# Installs the UNMODIFIED reference (dfm/celerite, /root/reference) into baseline/_ref, which is
# git-ignored but travels to the GPU box.  Used by tests/test_gpu_dropin.py, tests/test_terms.py,
# tests/test_reference_suite.py and tests/golden/make_golden*.py.  /root/reference is read-only,
# so the build runs from a copy under /tmp; the one addition is celerite/celerite_version.py,
# which setuptools_scm would have written (it is absent from this image).
# The reference's own test files are placed beside the install (baseline/_ref/ref_tests) so that
# they can be run, unmodified, against the replacement module on the GPU box, where
# /root/reference does not exist.  Nothing from the reference enters git history.
set -euo pipefail
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
REF="${REF:-/root/reference}"
DST="$ROOT/baseline/_ref"
if [ ! -d "$REF/celerite" ]; then
  echo "reference tree not mounted at $REF: keeping baseline/_ref as it is"; exit 0
fi
if [ -f "$DST/celerite/__init__.py" ] && ls "$DST"/celerite/solver*.so >/dev/null 2>&1 \
   && [ -f "$DST/ref_tests/test_celerite.py" ] && [ "${1:-}" != "--force" ]; then
  echo "up to date: $DST"; exit 0
fi
TMP="$(mktemp -d /tmp/celerite_ref.XXXXXX)"
trap 'rm -rf "$TMP"' EXIT
mkdir -p "$TMP/src"
# the Eigen tree holds ~10 MB of tests/docs the build does not need
(cd "$REF" && tar cf - --exclude=docs --exclude=paper --exclude=examples --exclude='cpp/lib/eigen/test' \
   --exclude='cpp/lib/eigen/doc' --exclude='cpp/lib/eigen/bench' --exclude='cpp/lib/eigen/unsupported/test' .) | (cd "$TMP/src" && tar xf -)
echo '__version__ = "0.4.3.dev0+reference"' > "$TMP/src/celerite/celerite_version.py"
rm -rf "$DST"; mkdir -p "$DST"
python -m pip install --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse \
  --target "$DST" "$TMP/src" 2>&1 | tail -3
mkdir -p "$DST/ref_tests"
cp "$REF/tests/test_celerite.py" "$REF/tests/test_terms.py" "$DST/ref_tests/"
echo "installed: $DST"
