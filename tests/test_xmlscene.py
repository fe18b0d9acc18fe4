"""CPU tests of the XML scene front end (beifong_b200/xmlscene.py): the reference's scene format drives the same
FlatScene as the programmatic configs.  Checked through the oracle (bit-identical profiles)."""
import os

import numpy as np
import pytest

from beifong_b200 import _native as N
from beifong_b200 import configs, xmlscene
from beifong_b200.scene import BeifongError
from oracle_lib import OracleScene

REF_XML = "/root/reference/python_scripts/trans_rad.xml"

# the scene of python_scripts/trans_rad.xml re-typed (same elements and values)
C1_XML = """
<scene version="2.1.0">
  <default name="spp" value="10000"/> <default name="resx" value="1"/> <default name="resy" value="1"/>
  <integrator type="time"><integrator type="pathtime"/></integrator>
  <shape type="rectangle" id="receiveAntenna">
    <transform name="to_world"><scale x="0.05" y="0.05"/><lookat origin="0, 0, 0" target="0, -1, 0" up="0, 0, 1"/></transform>
    <sensor type="fluxmeter">
      <film type="hdrfilm"><integer name="width" value="$resx"/><integer name="height" value="$resy"/><rfilter type="box"/></film>
      <sampler type="independent"><integer name="sample_count" value="$spp"/></sampler>
    </sensor>
  </shape>
  <bsdf type="twosided" id="material"><bsdf type="diffuse"><spectrum value="1.0" name="reflectance"/></bsdf></bsdf>
  <emitter type="spot">
    <spectrum value="1.0" name="intensity"/><float name="cutoff_angle" value="25"/><float name="beam_width" value="20"/>
    <transform name="to_world"><lookat origin="0, 0, 0" target="0, -1, 0" up="0, 0, 1"/></transform>
  </emitter>
  <shape type="rectangle" id="target">
    <transform name="to_world"><scale x="1" y="1"/><lookat origin="0, -4, 0" target="0, 0, 0" up="0, 0, 1"/></transform>
    <ref id="material" name="bsdf"/>
  </shape>
  <shape type="rectangle" id="gnd">
    <transform name="to_world"><scale x="20" y="20"/><lookat origin="0, 0, -0.5" target="0, 0, 0.5"/></transform>
    <ref id="material" name="bsdf"/>
  </shape>
</scene>
"""


def test_c1_xml_equals_programmatic_config():
    L = xmlscene.load_string(C1_XML)
    assert (L.sample_count, L.seed) == (10000, 0)
    assert L.shape_ids == {"receiveAntenna": 0, "target": 1, "gnd": 2}
    assert L.scene.profile_shape() == (1, 1, 155)
    a = OracleScene(L.scene.export(), accel=OracleScene.BRUTE).receive(20000, seed=0, threads=4)
    b = OracleScene(configs.c1_trans_rad().export(), accel=OracleScene.BRUTE).receive(20000, seed=0, threads=4)
    assert np.array_equal(a, b) and a[0, 0, 5:].sum() > 0


def test_param_override_and_undefined_param():
    L = xmlscene.load_string(C1_XML, spp=77)
    assert L.sample_count == 77
    with pytest.raises(BeifongError, match="undefined parameter"):
        xmlscene.load_string(C1_XML.replace('value="$spp"', 'value="$nope"'))


@pytest.mark.skipif(not os.path.exists(REF_XML), reason="reference tree not present")
def test_literal_reference_file_loads_identically():
    a = xmlscene.load_file(REF_XML).scene.export()
    b = xmlscene.load_string(C1_XML).scene.export()
    assert np.array_equal(a, b)


def test_unsupported_plugins_raise_like_the_reference():
    with pytest.raises(NotImplementedError):
        xmlscene.load_string(C1_XML.replace('type="diffuse"', 'type="plastic"'))
    with pytest.raises(NotImplementedError):
        xmlscene.load_string(C1_XML.replace('<shape type="rectangle" id="gnd">', '<shape type="sphere" id="gnd">'))
    with pytest.raises(NotImplementedError):
        xmlscene.load_string(C1_XML.replace('<rfilter type="box"/>', '<rfilter type="gaussian"/>'))
    with pytest.raises(BeifongError):
        xmlscene.load_string(C1_XML.replace('</scene>', '<shape type="ply"><string name="filename" value="nope.ply"/></shape></scene>'))


RADAR_XML = """
<scene version="2.1.0">
  <integrator type="phase"><integrator type="pathtimefrequency"><integer name="max_depth" value="3"/></integrator>
    <integer name="bins" value="2"/></integrator>
  <shape type="rectangle" id="rx">
    <transform name="to_world"><scale x="0.0025" y="0.0025"/><rotate y="1" angle="90"/><translate x="0" y="0" z="0.3"/></transform>
    <receiver type="wignerreceiver">
      <string name="signaltype" value="linfmcw"/><float name="crf" value="5000"/><float name="chirp_len" value="200e-6"/>
      <float name="freq_centre" value="94e9"/><float name="freq_sweep" value="6e9"/>
      <float name="adc_sampling_start" value="20e-6"/><float name="adc_sampling_end" value="180e-6"/>
      <string name="receive_type" value="mix_resample"/>
      <spectrum name="antenna_texture" value="1.0"/>
      <adc type="hdradc"><integer name="t_bins" value="1"/><integer name="f_bins" value="256"/>
        <float name="t_bandwidth" value="200e-6"/><float name="f_bandwidth" value="6e6"/><rfilter type="box"/></adc>
      <sampler type="independent"><integer name="sample_count" value="4096"/><integer name="seed" value="3"/></sampler>
    </receiver>
  </shape>
  <shape type="rectangle" id="tx">
    <transform name="to_world"><scale x="0.0025" y="0.0025"/><rotate y="1" angle="90"/><translate x="0" y="0.01" z="0.3"/></transform>
    <transmitter type="wignertransmitter">
      <string name="signaltype" value="linfmcw"/><float name="crf" value="5000"/><float name="chirp_len" value="200e-6"/>
      <float name="freq_centre" value="94e9"/><float name="freq_sweep" value="6e9"/><boolean name="resample_freq" value="true"/>
      <spectrum name="antenna_texture" value="1.0"/>
    </transmitter>
  </shape>
  <shape type="rectangle" id="wall">
    <transform name="to_world"><scale x="5" y="5"/><rotate y="1" angle="-90"/><translate x="6" y="0" z="0"/></transform>
    <bsdf type="twosided"><bsdf type="roughconductor"><float name="alpha" value="0.3"/><string name="distribution" value="ggx"/></bsdf></bsdf>
  </shape>
</scene>
"""


def test_gen3_radar_scene_from_xml():
    from beifong_b200.scene import RADAR_94GHZ
    L = xmlscene.load_string(RADAR_XML, consts=RADAR_94GHZ)
    assert (L.sample_count, L.seed) == (4096, 3)
    assert L.scene.profile_shape() == (256, 1, 5)            # Y, A, W + 2 phase bins
    prof = OracleScene(L.scene.export(), accel=OracleScene.BRUTE).receive(100000, seed=3, threads=4)
    k = int(6.0 / (6e6 / 256 / (6e9 / 200e-6) * 299792458.0 / 2))       # wall at 6 m -> beat-frequency bin 51
    assert prof[:k - 1, 0, 0].sum() == 0 and prof[k - 1:k + 6, 0, 0].sum() > 0.5 * prof[..., 0].sum()
    assert prof[..., 3].sum() > 0 and prof[..., 4].sum() > 0           # both phase bins populated


ARRAY_XML = """<integer name="n_elems" value="4"/><vector name="steering_vector" x="0" y="0.05" z="0"/>
      <vector name="elem_dims" x="0.005" y="0.00125" z="1"/><vector name="elem_spacing" x="0.00125" y="0.00125" z="0.00125"/>
      <vector name="elem_axis" x="0" y="1" z="0"/>
      <transform name="array_loc"><rotate y="1" angle="90"/><translate x="0" y="%s" z="0.3"/></transform>"""


def test_phased_endpoints_from_xml():
    """phasedtransmitter / phasedreceiver keep the reference's parameter names (phasedtransmitter.cpp:104-117)."""
    from beifong_b200.scene import RADAR_94GHZ
    from beifong_b200 import _native as N
    text = (RADAR_XML.replace('<receiver type="wignerreceiver">', '<receiver type="phasedreceiver">' + ARRAY_XML % "0")
            .replace('<transmitter type="wignertransmitter">', '<transmitter type="phasedtransmitter">' + ARRAY_XML % "0.01"))
    L = xmlscene.load_string(text, consts=RADAR_94GHZ)
    orc = OracleScene(L.scene.export(), accel=OracleScene.BRUTE)
    prof = orc.receive(50000, seed=3, threads=4)
    assert np.all(np.isfinite(prof)) and np.abs(prof[..., 0]).sum() > 0
    # the same array built programmatically gives the same flat scene
    ref = xmlscene.load_string(RADAR_XML, consts=RADAR_94GHZ)
    assert len(L.scene.export()) == len(ref.scene.export())
    with pytest.raises(BeifongError):     # to_world is not allowed on the transmitter itself (phasedtransmitter.cpp:54-57)
        xmlscene.load_string(text.replace('<transmitter type="phasedtransmitter">',
                                          '<transmitter type="phasedtransmitter"><transform name="to_world"><translate x="1"/></transform>'),
                             consts=RADAR_94GHZ)
    with pytest.raises(BeifongError):     # n_elems outside [1, 64]
        xmlscene.load_string(text.replace('<integer name="n_elems" value="4"/>', '<integer name="n_elems" value="80"/>'), consts=RADAR_94GHZ)


def test_adc_reconstruction_filters_from_xml():
    """<rfilter> under an <adc>: gaussian (stddev -> radius 4 stddev), tent, and the gaussian default when absent
    (adc.cpp:60-76); the oracle then takes the wide branch of SignalBlock::put and spreads the wall's peak."""
    from beifong_b200.scene import RADAR_94GHZ
    box = OracleScene(xmlscene.load_string(RADAR_XML, consts=RADAR_94GHZ).scene.export(), accel=OracleScene.BRUTE).receive(20000, seed=3, threads=4)
    for rf in ('<rfilter type="gaussian"><float name="stddev" value="0.5"/></rfilter>', '<rfilter type="tent"/>', '',
               '<rfilter type="mitchell"/>', '<rfilter type="catmullrom"/>', '<rfilter type="lanczos"><integer name="lobes" value="2"/></rfilter>'):
        L = xmlscene.load_string(RADAR_XML.replace('<rfilter type="box"/>', rf), consts=RADAR_94GHZ)
        wide = OracleScene(L.scene.export(), accel=OracleScene.BRUTE).receive(20000, seed=3, threads=4)
        assert wide.shape == box.shape and np.count_nonzero(wide[..., 0]) > np.count_nonzero(box[..., 0])
    with pytest.raises(NotImplementedError):
        xmlscene.load_string(RADAR_XML.replace('<rfilter type="box"/>', '<rfilter type="mitchell"><float name="B" value="0.5"/></rfilter>'), consts=RADAR_94GHZ)


def test_include_alias_and_serialized_shapes(tmp_path):
    """<include> (xml.cpp:653-705), <alias> (xml.cpp:594-613) and <shape type="serialized"> (serialized.cpp) build the same
    scene as the inline form."""
    from beifong_b200 import meshes
    from beifong_b200.scene import RADAR_94GHZ
    p, f = meshes.car_body(12, 7)
    meshes.save_serialized(str(tmp_path / "body.serialized"), [(meshes.stairs(3)[0], meshes.stairs(3)[1], None), (p, f, None)])
    (tmp_path / "materials.xml").write_text(
        '<scene version="2.1.0"><bsdf type="twosided" id="metal"><bsdf type="roughconductor"><float name="alpha" value="0.3"/>'
        '<string name="distribution" value="ggx"/></bsdf></bsdf><alias id="metal" as="hull"/></scene>')
    (tmp_path / "target.xml").write_text(
        '<shape type="serialized" id="car"><string name="filename" value="body.serialized"/><integer name="shape_index" value="1"/>'
        '<transform name="to_world"><translate x="5" y="0" z="0"/></transform><ref id="hull"/></shape>')
    text = RADAR_XML.replace("</scene>", '<include filename="materials.xml"/><include filename="target.xml"/><alias id="car" as="target"/></scene>')
    L = xmlscene.load_string(text, base_dir=str(tmp_path), consts=RADAR_94GHZ)
    assert L.shape_ids["target"] == L.shape_ids["car"]
    inline = RADAR_XML.replace("</scene>", '<bsdf type="twosided" id="metal"><bsdf type="roughconductor"><float name="alpha" value="0.3"/>'
                               '<string name="distribution" value="ggx"/></bsdf></bsdf>'
                               '<shape type="serialized" id="car"><string name="filename" value="body.serialized"/><integer name="shape_index" value="1"/>'
                               '<transform name="to_world"><translate x="5" y="0" z="0"/></transform><ref id="metal"/></shape></scene>')
    L2 = xmlscene.load_string(inline, base_dir=str(tmp_path), consts=RADAR_94GHZ)
    assert np.array_equal(L.scene.export(), L2.scene.export())
    with pytest.raises(BeifongError, match="not found"):
        xmlscene.load_string(RADAR_XML.replace("</scene>", '<include filename="nope.xml"/></scene>'), base_dir=str(tmp_path), consts=RADAR_94GHZ)
    with pytest.raises(BeifongError, match="referenced id"):
        xmlscene.load_string(RADAR_XML.replace("</scene>", '<alias id="ghost" as="x"/></scene>'), consts=RADAR_94GHZ)
    (tmp_path / "loop.xml").write_text('<scene version="2.1.0"><include filename="loop.xml"/></scene>')
    with pytest.raises(BeifongError, match="recursion limit"):
        xmlscene.load_string(RADAR_XML.replace("</scene>", '<include filename="loop.xml"/></scene>'), base_dir=str(tmp_path), consts=RADAR_94GHZ)
