"""Seeded scenes shared by the parity tests (sizes the CPU oracle finishes in seconds)."""
import dynosam_b200 as d


def simulator_cfg1(n_frames=50):
    """BASELINE config 1: the reference's makeDefaultScenario(), 2 objects, 50 frames."""
    return d.Scene.simulator(n_frames)


def small_noisy(seed=7, n_frames=40, n_objects=3, static_per_frame=50, dynamic_per_object=12):
    return d.Scene.synthetic(n_frames=n_frames, n_objects=n_objects, static_per_frame=static_per_frame,
                             dynamic_per_object=dynamic_per_object, static_track_max=20, dynamic_track_max=15, seed=seed)


def medium_noisy(seed=11):
    return d.Scene.synthetic(n_frames=120, n_objects=4, static_per_frame=120, dynamic_per_object=25, seed=seed)
