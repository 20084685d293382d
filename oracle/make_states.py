"""TEST/BENCH INPUT GENERATOR (runs in the build container, uses the oracle + SciPy).

Writes deepflame-dev_b200/data/states_<mech>.npz: partially-burnt states sampled along
constant-pressure adiabatic ignition trajectories (the df0DFoam situation), used by bench.py and
the tests as the "burning" half of the synthetic cell batches.
usage: python -m oracle.make_states <mechanism.yaml> <fuel> <out.npz> [n_per_traj]
"""
import sys
import numpy as np
from . import mech, states


def main():
    path, fuel, out = sys.argv[1:4]
    n = int(sys.argv[4]) if len(sys.argv) > 4 else 256
    m = mech.load_mechanism(path)
    Ts, Ys, tag = [], [], []
    for (T0, p_atm, phi, t_end) in [(1200.0, 1.0, 1.0, 2e-2), (1500.0, 5.0, 0.7, 2e-3),
                                    (1800.0, 10.0, 1.5, 2e-4), (1400.0, 2.0, 1.0, 5e-3)]:
        Y0 = states.fuel_air(m, fuel, [phi])[0]
        T, Y = states.ignition_trajectory(m, T0, p_atm * 101325.0, Y0, t_end, n)
        Ts.append(T); Ys.append(Y); tag += [(T0, p_atm, phi)] * len(T)
        print(T0, p_atm, phi, "->", T[0], T[-1], flush=True)
    np.savez_compressed(out, species=np.array(m.species), T=np.concatenate(Ts), Y=np.concatenate(Ys),
                        origin=np.array(tag))


if __name__ == "__main__":
    main()
