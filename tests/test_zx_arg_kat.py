"""The reference's ARGMIN / ARGMAX known-answer tests (argmin_tests.cpp:22-107, argmax_tests.cpp:22-105; vectors in
tests/golden/reference_kat_arg.json, every fixed-width value type but bool) through the CUDA path."""
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
if not torch.cuda.is_available():  # pragma: no cover
    pytest.skip("needs a CUDA device", allow_module_level=True)

from tests import gpu_util, kat  # noqa: E402

CASES = kat.load_cases("reference_kat_arg.json")


@pytest.mark.parametrize("name,c,kt,vt", CASES, ids=[x[0] for x in CASES])
def test_reference_argmin_argmax_known_answers(name, c, kt, vt):
    keys, vals, exp = kat.materialize(c, kt, vt)
    G, kout, aout, _ = gpu_util.run_gpu([keys], [vals], [(0, c["agg"], c["ddof"])], include_null_keys=c["include_null_keys"])
    kat.check(name, G, (kout[0][0], kout[0][1], None), (aout[0][0], aout[0][1], None), exp)
