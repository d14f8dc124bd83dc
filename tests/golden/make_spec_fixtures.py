"""Regenerates tests/golden/reference_specs.json from the reference checkout (run in the build container,
where /root/reference exists; the GPU box only sees the committed JSON).

Sources: the serialised query specs the reference documents and tests
  docs/benchmark/druid/queries/*.json                               (8 GroupBy specs as Utils.jsonFormat renders them)
  src/test/scala/.../client/test/SparklineSQLTest.scala:73-280      (raw specs sent through EXECUTE QUERY)
Each entry keeps the spec verbatim (minified) so the engine's parser is tested on the reference's own bytes."""
import glob
import json
import os
import re

REF = "/root/reference"
out = {}
for p in sorted(glob.glob(os.path.join(REF, "docs/benchmark/druid/queries/*.json"))):
    out["docs/" + os.path.basename(p)[:-5]] = json.load(open(p))
src = open(os.path.join(REF, "src/test/scala/org/sparklinedata/druid/client/test/SparklineSQLTest.scala")).read()
for i, m in enumerate(re.finditer(r'test\("([^"]+)"\).*?execute query\s*\n(.*?)"""', src, re.S)):
    body = "\n".join(l.split("|", 1)[1] if "|" in l else l for l in m.group(2).splitlines())
    try:
        out["sqltest/" + m.group(1)] = json.loads(body)
    except json.JSONDecodeError:
        pass
json.dump(out, open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "reference_specs.json"), "w"), separators=(",", ":"), sort_keys=True)
print(len(out), "specs:", ", ".join(out))
