"""Constants and helper functions of the reference's ``set_parameters.m`` that the
hot path consumes.  Same names, same values, same
operation order; file/directory handles of the script are out of scope.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np


@dataclass(frozen=True)
class Parameters:
    # physical constants (set_parameters.m:5-8)
    lya_wavelength: float = 1215.6701
    lyb_wavelength: float = 1025.7223
    lyman_limit: float = 911.7633
    speed_of_light: float = 299792458.0
    # preprocessing (set_parameters.m:25-26)
    z_qso_cut: float = 2.15
    min_num_pixels: int = 200
    # null model (set_parameters.m:33-36)
    min_lambda: float = 911.75
    max_lambda: float = 1215.75
    dlambda: float = 0.25
    k: int = 20
    # optimisation initial values (set_parameters.m:40-42)
    initial_c_0: float = 0.1
    initial_tau_0: float = 0.0023
    initial_beta: float = 3.65
    # DLA samples (set_parameters.m:48-53)
    num_dla_samples: int = 10000
    alpha: float = 0.9
    uniform_min_log_nhi: float = 20.0
    uniform_max_log_nhi: float = 23.0
    fit_min_log_nhi: float = 20.0
    fit_max_log_nhi: float = 22.0
    # instrumental broadening (set_parameters.m:59-60)
    width: int = 3
    pixel_spacing: float = 1e-4
    # absorber model (set_parameters.m:63)
    num_lines: int = 3

    # set_parameters.m:11
    def kms_to_z(self, kms):
        return (kms * 1000) / self.speed_of_light

    # set_parameters.m:14-18
    def emitted_wavelengths(self, observed_wavelengths, z):
        return observed_wavelengths / (1 + z)

    def observed_wavelengths(self, emitted_wavelengths, z):
        return emitted_wavelengths * (1 + z)

    @property
    def prior_z_qso_increase(self):  # set_parameters.m:56
        return self.kms_to_z(30000)

    @property
    def max_z_cut(self):  # set_parameters.m:65
        return self.kms_to_z(3000)

    @property
    def min_z_cut(self):  # set_parameters.m:69
        return self.kms_to_z(3000)

    # set_parameters.m:66-67
    def max_z_dla(self, wavelengths, z_qso):
        return (np.max(wavelengths) / self.lya_wavelength - 1) - self.max_z_cut

    # set_parameters.m:70-73
    def min_z_dla(self, wavelengths, z_qso):
        return max(np.min(wavelengths) / self.lya_wavelength - 1,
                   self.observed_wavelengths(self.lyman_limit, z_qso) / self.lya_wavelength - 1 + self.min_z_cut)


DEFAULT = Parameters()
