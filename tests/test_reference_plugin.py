"""The product as a plugin of the reference: eradiate-kernel_b200/plugin/volpath_b200.cpp is compiled against the
reference's headers (oracle/Makefile.ref) and loaded by the UNMODIFIED reference binary's PluginManager when a
scene says <integrator type="volpath_b200">.  Everything around it -- XML loader, scene, sensors, film -- is
the reference's own code; Integrator::render reads that object graph and calls the C-ABI.

* CPU: the plugin loads and fails loudly without a device; its describe() is checked without a GPU by rendering
  the descriptor it produces with the oracle (refprobe --describe-oracle) against the reference's own films.
* GPU: the reference binary renders the fixture scenes with the plugin on the B200; films against the films of
  its own CPU volpath (tests/golden/ref_golden.npz).

Needs oracle/_ref (built where /root/reference exists; it travels to the GPU box); skipped elsewhere."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import make_ref_golden  # noqa: E402
from test_reference_parity import REF, check_against_reference  # noqa: E402

# every fixture scene; a `moment` wrapper becomes the plugin's own moment=true (the wrapper calls sample(), which the
# GPU path does not have)
PLUGIN_CASES = sorted(make_ref_golden.CASES)


def ref_xml_or_skip():
    from oracle import ref_xml
    if not (ref_xml.available() and os.path.exists(ref_xml.PLUGIN)):
        pytest.skip("oracle/_ref (reference binary + volpath_b200 plugin) is not built on this machine")
    return ref_xml


def test_plugin_fails_loudly_without_a_device(pkg):
    ref_xml = ref_xml_or_skip()
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    d, variant = make_ref_golden.scene_dict("c1")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ref_xml.render(ref_xml.with_plugin(d), variant, write_vol=pkg.scene.write_vol)


@pytest.mark.parametrize("name", PLUGIN_CASES)
def test_plugin_describe_against_reference_films(pkg, oracle, name):
    """describe() on the reference's object graph -> B2vpSceneDesc -> CPU oracle == the reference's own volpath."""
    ref_xml = ref_xml_or_skip()
    d, variant = make_ref_golden.scene_dict(name)
    film, _ = ref_xml.render(ref_xml.with_plugin(d), variant, write_vol=pkg.scene.write_vol, describe_oracle=True)
    check_against_reference(name, film, REF[name], variant)
    if variant == "scalar_mono_double":
        # the Python host mirror (load_dict) and the C++ plugin describe the scene identically: same bits
        r = oracle.render(pkg, pkg.load_dict(d, variant), pkg.abi.SEED_SCALAR_BLOCKS)
        assert np.array_equal(r.film, film), name


def test_plugin_rejects_content_outside_the_scope(pkg):
    ref_xml = ref_xml_or_skip()
    d, variant = make_ref_golden.scene_dict("c2_paths")
    d["sensor"]["film"]["rfilter"] = {"type": "gaussian"}   # a reconstruction filter wider than a pixel
    with pytest.raises(RuntimeError, match="outside the scope"):
        ref_xml.render(ref_xml.with_plugin(d), variant, write_vol=pkg.scene.write_vol, describe_oracle=True)


@pytest.mark.gpu
@pytest.mark.parametrize("name", PLUGIN_CASES)
def test_reference_binary_renders_on_the_gpu_through_the_plugin(pkg, name):
    ref_xml = ref_xml_or_skip()
    d, variant = make_ref_golden.scene_dict(name)
    film, seconds = ref_xml.render(ref_xml.with_plugin(d), variant, write_vol=pkg.scene.write_vol)
    check_against_reference(name, film, REF[name], variant)
    assert seconds > 0
