"""Physical constants -- same names and units as the reference's ``Model/constants.py:1-19``.

The reference pulls twelve of these from the third-party ``sympl`` package
(``sympl.get_constant``); ``sympl`` is an unpinned dependency that is not part of the
reference tree.  The values below are sympl 0.4.x ``default_constants``.  Only
``p_surface_earth`` is pinned by a reference artefact (``CO2_no_sw.npy`` has
``p.max() == 101320.00000000007`` and ``real_gas.py:241`` documents 101320 Pa); the others
are "parity unpinned" -- the oracle shim (``oracle/reference_shim.py``) feeds the *same*
numbers to the unmodified reference, so every parity test compares like with like.

Units: SI; wavenumbers in cm^-1 elsewhere in the package.
"""

g = 9.80665                      # gravitational_acceleration, m/s^2
c_p_dry = 1004.64                # heat_capacity_of_dry_air_at_constant_pressure, J/kg/K
sigma = 5.670367e-8              # stefan_boltzmann_constant, W/m^2/K^4
p_surface_earth = 1.0132e5       # reference_air_pressure, Pa
p_one_atmosphere = 101325        # one atmosphere in Pa
p_toa_earth = 20                 # top of atmosphere, Pa
F_sun = 1367.0                   # solar_constant, W/m^2
Omega = 7.292e-5                 # planetary_rotation_rate, s^-1
R_earth = 6.37122e6              # planetary_radius, m
R_specific = 287.0               # gas_constant_of_dry_air, J/kg/K
Avogadro = 6.022140857e23        # avogadro_constant, mole^-1
speed_of_light = 299792458.0     # m/s
h_planck = 6.62607004e-34        # J s
k_boltzmann = 1.38064852e-23     # J/K
AU = 1.495978707e11              # average earth-sun distance, m
R_sun = 6.96340e8                # radius of sun, m
T_sun = 5778                     # temperature of sun, K
