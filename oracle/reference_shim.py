"""TEST INFRASTRUCTURE ONLY -- never imported by the product package.

Import shim that lets the UNMODIFIED reference (``/root/reference``) run in this
container so that golden vectors can be generated from the reference itself.

The reference needs two modules that are not installed here and cannot be
installed (no network):

* ``sympl``      -- only ``sympl.get_constant`` is used (Model/constants.py:1-16)
* ``matplotlib`` -- imported at module top for plotting only
                   (Model/radiation/real_gas.py:7, base.py has none, grey.py:9,
                   real_gas_data/hitran.py:7, animation.py)

Both are replaced by stubs registered in ``sys.modules``.  ``np.trapz`` is
aliased to ``np.trapezoid`` when a future NumPy drops it.

``/root/reference`` exists only in the build container.  On the GPU box the shim
falls back to the unmodified copy staged under ``oracle/_ref`` (oracle/make_ref.py):
that is what ``bench.py --impl reference`` and the ``cpu_baseline`` leg time.  The
``-m gpu`` tests and ``smoke()`` never import this file.
"""
import importlib
import os
import sys
import types
import warnings

import numpy as np

# The reference itself in the build container; on the GPU box (no /root/reference) the unmodified copy
# that oracle/make_ref.py staged under oracle/_ref (git-ignored, travels with the snapshot).
_STAGED = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")
REFERENCE_ROOT = os.environ.get("CLIMATEMODEL_REFERENCE_ROOT") or (
    "/root/reference" if os.path.isdir("/root/reference/Model/radiation") else _STAGED)

# sympl 0.4.x default_constants values for the 12 names Model/constants.py:3-16 asks for.
# Only reference_air_pressure is pinned by a reference artefact (CO2_no_sw.npy p.max()=101320);
# the rest are "parity unpinned" third-party values -- harmless because the oracle and the
# CUDA path are fed the same numbers (climatemodel_b200/constants.py holds the same table).
SYMPL_CONSTANTS = {
    "gravitational_acceleration": 9.80665,
    "heat_capacity_of_dry_air_at_constant_pressure": 1004.64,
    "stefan_boltzmann_constant": 5.670367e-8,
    "reference_air_pressure": 1.0132e5,
    "solar_constant": 1367.0,
    "planetary_rotation_rate": 7.292e-5,
    "planetary_radius": 6.37122e6,
    "gas_constant_of_dry_air": 287.0,
    "avogadro_constant": 6.022140857e23,
    "speed_of_light": 299792458.0,
    "planck_constant": 6.62607004e-34,
    "boltzmann_constant": 1.38064852e-23,
}


def reference_available():
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "Model", "radiation"))


def _install_stubs():
    if "sympl" not in sys.modules:
        sympl = types.ModuleType("sympl")

        def get_constant(name, units=None):
            return SYMPL_CONSTANTS[name]

        sympl.get_constant = get_constant
        sys.modules["sympl"] = sympl
    try:
        import matplotlib  # noqa: F401
    except ImportError:
        mpl = types.ModuleType("matplotlib")
        mpl.use = lambda *a, **k: None
        mpl.__path__ = []
        pyplot = types.ModuleType("matplotlib.pyplot")
        animation = types.ModuleType("matplotlib.animation")
        animation.FuncAnimation = object
        widgets = types.ModuleType("matplotlib.widgets")
        widgets.Slider = object
        widgets.Button = object
        mpl.pyplot, mpl.animation, mpl.widgets = pyplot, animation, widgets
        sys.modules.update({"matplotlib": mpl, "matplotlib.pyplot": pyplot,
                            "matplotlib.animation": animation, "matplotlib.widgets": widgets})
        tk = types.ModuleType("mpl_toolkits")
        tk.__path__ = []
        ag = types.ModuleType("mpl_toolkits.axes_grid1")
        ag.make_axes_locatable = lambda *a, **k: None
        tk.axes_grid1 = ag
        sys.modules.update({"mpl_toolkits": tk, "mpl_toolkits.axes_grid1": ag})
    if not hasattr(np, "trapz"):
        np.trapz = np.trapezoid


def load_reference():
    """Return the reference's ``Model.radiation.real_gas`` module (imported unmodified)."""
    if not reference_available():
        raise RuntimeError("reference tree not present at %s" % REFERENCE_ROOT)
    _install_stubs()
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    warnings.filterwarnings("ignore", category=DeprecationWarning)
    warnings.filterwarnings("ignore", category=RuntimeWarning)
    return importlib.import_module("Model.radiation.real_gas")


def set_lookup_folder(folder):
    """Point the reference at another LookupTables folder (hitran.py:24; real_gas.py:1 binds by name)."""
    rg = load_reference()
    hitran = importlib.import_module("Model.radiation.real_gas_data.hitran")
    if not folder.endswith("/"):
        folder = folder + "/"
    hitran.LookupTableFolder = folder
    rg.LookupTableFolder = folder
    return folder


def default_lookup_folder():
    return os.path.join(REFERENCE_ROOT, "Model/radiation/real_gas_data/HitranData/LookupTables/")
