#!/bin/bash
# One batched GPU test command: every @pytest.mark.gpu test, JUnit report under gpurun_out/.
#   /usr/local/graft/bin/gpurun --timeout 900 -- 'bash tools/run_gpu_tests.sh'
mkdir -p gpurun_out
timeout 800 python -m pytest tests -m gpu -q --junitxml=gpurun_out/gpu_tests.xml 2>&1 | tail -5
python - <<'PY'
import xml.etree.ElementTree as ET
r = ET.parse("gpurun_out/gpu_tests.xml").getroot()
s = r if r.tag == "testsuite" else r.find("testsuite")
print("junit:", {k: s.get(k) for k in ("tests", "failures", "errors", "skipped", "time")})
PY
