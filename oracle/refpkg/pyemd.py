"""Stand-in for pyemd 0.5.1 (third-party, absent from the reference tree and from this image).
The EMD value of side_effect_score is therefore "parity unpinned" (DESIGN.md).  The oracle harness
patches safelife.side_effects.earth_mover_distance to capture the pre-EMD density maps instead."""


def emd(first_histogram, second_histogram, distance_matrix, extra_mass_penalty=-1.0):
    raise NotImplementedError("pyemd is not installed; EMD parity is unpinned")
