#!/bin/bash
# tools/run_asserts.sh -- on the GPU box: the ray-level, image and front-half tests against the debug-assert build kept under ab/asserts
# (make EXTRA=-DPATHS_DEBUG_ASSERTS; tools/build_variant.sh asserts -DPATHS_DEBUG_ASSERTS), then the counters it keeps.
# compute-sanitizer is closed on this pool; this run is what stands in for it.
set -o pipefail
cd "$(dirname "$0")/.."
export PATHS_B200_LIB=$PWD/ab/asserts/paths_b200/lib/libpaths_b200.so
python - <<'PY'
import paths_b200 as pb
print("library:", pb.lib().paths_version().decode(), "from", pb._capi.LIB_PATH)
assert b"+asserts" in pb.lib().paths_version()
PY
python tools/kernel_families_smoke.py 2>&1 | tail -3
python -m pytest tests/test_gpu_production_rays.py tests/test_gpu_parity.py tests/test_gpu_extensions.py tests/test_gpu_frontend.py tests/test_gpu_edges.py tests/test_zz_no_alloc.py -q -m gpu -p no:cacheprovider 2>&1 | tail -6
python - <<'PY'
import paths_b200 as pb
for name, prec in (("teapot", 32), ("bunny", 32), ("dragon", 32), ("bunny", 48)):
    import sys; sys.path.insert(0, "tools"); import make_assets; make_assets.ensure("dragon")
    sc = pb.Scene.load(f"scenes/{name}.yml", width=1920, height=1080)
    r = pb.Renderer(sc.camera, sc, precision=prec, seed=9)
    r.render_samples(32 if prec == 32 else 4)
    st = r.stats()
    r.shutdown()
    print(f"{name} precision {prec}: samples {st['samples']} extend_rays {st['extend_rays']} shadow_rays {st['shadow_rays']} "
          f"debug_asserts {st['debug_asserts']} stack_overflows {st['stack_overflows']} assert_failures {st['assert_failures']} bad_samples {st['bad_samples']}")
    assert st["debug_asserts"] == 1 and st["stack_overflows"] == 0 and st["assert_failures"] == 0
print("debug-assert build: all counters zero")
PY
