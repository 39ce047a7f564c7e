"""CPU tier: the fork's compound cart (reference models/double_cart_1.urdf:15-80, used by random_action_agent.py:781-794)
as the composite rigid body with per-face friction that urdf.parse_compound_cart_urdf builds: constants against the
reference's files (tests/golden/double_cart_constants.json), and what the per-face friction means physically - the cart
slides freely on the ground while it drags the pole along - on the oracle and on the kernels' host program."""
import json
import os

import numpy as np

from cartpoleplusplus_b200 import config as C
from cartpoleplusplus_b200 import default_opts, make_config, urdf
from hostharness import HostKernels
from oracle import Oracle

GOLD = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "double_cart_constants.json")))


def test_packaged_double_cart_restates_the_reference_files():
    import xml.etree.ElementTree as ET
    root = ET.parse(os.path.join(urdf.MODEL_DIR, "double_cart.urdf")).getroot()
    for ref_name in ("double_cart_1", "double_cart_2"):   # the two files differ in names and colours only
        ref = GOLD[ref_name]
        assert ref["joint"]["type"] == root.find("joint").attrib["type"] == "fixed"
        mine = {l.attrib["name"]: l for l in root.findall("link")}
        for (rn, rl), key in zip(sorted(ref["links"].items(), key=lambda kv: -kv[1]["com"][2]), ("upperLinkWithFriction", "lowerFrictionlessLink")):
            l = mine[key]
            assert float(l.find("inertial/mass").attrib["value"]) == rl["mass"]
            assert [float(v) for v in l.find("inertial/origin").attrib["xyz"].split()] == rl["com"] == rl["box_origin"]
            assert [float(v) for v in l.find("collision/geometry/box").attrib["size"].split()] == rl["box"]
            assert float(l.find("contact/lateral_friction").attrib["value"]) == rl["contact"]["lateral_friction"]
            assert float(l.find("contact/inertia_scaling").attrib["value"]) == rl["contact"]["inertia_scaling"]
            ia = l.find("inertial/inertia").attrib
            assert np.allclose([float(ia[k]) for k in ("ixx", "iyy", "izz")], rl["inertia"], rtol=1e-9)


def test_composite_body():
    cart = urdf.parse_compound_cart_urdf(os.path.join(urdf.MODEL_DIR, "double_cart.urdf"))
    assert cart.mass == 1.0 and cart.box_size == (0.2, 0.2, 0.05)
    assert cart.lateral_friction == 1.0 and cart.ground_friction == 0.0
    # per half: 3 x box inertia about its own centre + m d^2 (d = 1.25 cm, untouched by the scaling)
    half = 0.5 / 12 * (0.2 ** 2 + 0.025 ** 2)
    want = (2 * (3 * half + 0.5 * 0.0125 ** 2),) * 2 + (2 * 3 * 0.5 / 12 * (0.2 ** 2 + 0.2 ** 2),)
    assert np.allclose(cart.effective_inertia(), want, rtol=1e-12)
    # without the scaling the composite is the single 5 cm box exactly
    one = urdf.parse_box_urdf(os.path.join(urdf.MODEL_DIR, "cart.urdf"))
    assert np.allclose([2 * (half + 0.5 * 0.0125 ** 2)], [one.box_inertia()[0]], rtol=1e-12)
    cfg = make_config(default_opts(cart_model="double_cart"), False, num_envs=1, dtype="fp64")
    assert cfg.cart_ground_friction == 0.0 and cfg.cart.friction == 1.0 and cfg.cart.mass == 1.0
    assert make_config(default_opts(), False, num_envs=1).cart_ground_friction == -1.0


def _push(cart_kw, backend):
    cfg = make_config(default_opts(initial_force=0.0, action_force=20.0, no_termination=True, **cart_kw), False, num_envs=1, dtype="fp64", seed=1)
    env = backend(cfg)
    env.reset()
    for _ in range(6):  # 12 substeps of a 20 N push along x
        env.step(np.array([[1.0, 0.0]], dtype=np.float32))
    return env.state[0].copy()


def test_slides_on_the_ground_and_drags_the_pole():
    for backend in (lambda c: Oracle(c, dtype="fp64"), lambda c: HostKernels(c, "fp64", fast=0)):
        compound = _push(dict(cart_model="double_cart"), backend)
        single = _push(dict(cart_friction=1.0), backend)
        # 20 N on 1 kg + 5 kg coupled by the friction under the pole's foot: were they one body, 3.3 m/s^2 for 11 substeps (the
        # force lags one substep) = 0.15 m/s; the foot is dragged and the pole leans back, so its centre lags and the cart leads
        vx_cart, vx_pole = compound[C.S_CART + 7], compound[C.S_POLE + 7]
        assert 0.2 < vx_cart < 0.35 and 0.08 < vx_pole < 0.16, (vx_cart, vx_pole)
        assert abs((1.0 * vx_cart + 5.0 * vx_pole) / 6.0 - 20.0 / 6.0 * 11 / 240.0) < 0.01   # momentum: nothing was lost to the ground
        # the single box with friction 1.0 also rubs on the ground (0.5 x 1.0 x 6 kg x g = 29 N > 20 N): it stays put
        assert abs(single[C.S_CART + 7]) < 1e-3
        # the fork's shipped cart.urdf (friction 0 everywhere): the cart shoots out from under the pole
        bare = _push(dict(), backend)
        assert bare[C.S_CART + 7] > 0.7 and abs(bare[C.S_POLE + 7]) < 2e-3


def test_oracle_and_kernel_program_agree_bit_for_bit_on_the_compound_cart():
    n = 24
    cfg = make_config(default_opts(initial_force=55.0, action_force=100.0, cart_model="double_cart", steps_per_repeat=3), False, num_envs=n,
                      dtype="fp64", seed=2)
    ora, par = Oracle(cfg, dtype="fp64"), HostKernels(cfg, "fp64", fast=0)
    assert np.array_equal(ora.reset(), par.reset())
    rng = np.random.default_rng(0)
    for t in range(15):
        a = rng.uniform(-1, 1, size=(n, 2)).astype(np.float32)
        assert all(np.array_equal(x, y) for x, y in zip(ora.step(a), par.step(a))), t
    assert np.array_equal(ora.state, par.state)
