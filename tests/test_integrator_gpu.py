"""GPU parity tests for the integrator extension (velocity, mass, dt, energy --
the reference has positions only, so the checker is oracle_nbody_integrate, itself pinned to the
reference's accel() through the unit-mass case in tests/test_oracle_cpu.py).
Positions and velocities: bit-identical to the oracle (the asked tolerances -- positions 1e-10 after
10 steps, energy 1e-12 relative -- are asserted too, so the numbers appear in the test)."""
import numpy as np
import pytest

from hg_c_benchmarks_b200 import synth

pytestmark = pytest.mark.gpu


def _case(n, seed):
    return synth.plummer(n, seed=seed), synth.velocities(n, seed=seed + 1), synth.masses(n, seed=seed + 2)


@pytest.mark.parametrize("n,steps,dt", [(1, 3, 0.01), (3, 10, 0.001), (100, 5, 0.01), (513, 4, 1e-3),
                                        (1024, 2, 0.125), (2000, 10, 1e-3), (4096, 10, 1e-3)])
def test_integrate_is_bit_identical_to_the_oracle(hgb, n, steps, dt):
    from oracle import port
    pos, vel, mass = _case(n, seed=n)
    wp, wv = pos.copy(), vel.copy()
    port.nbody_integrate(wp, wv, mass, dt, steps)
    with hgb.nbody.NBody(n) as nb:
        nb.upload(pos).set_dynamics(vel, mass, dt)
        nb.integrate(steps)
        gp, gv = nb.download(), nb.velocities()
    assert np.array_equal(gp, wp) and np.array_equal(gv, wv)
    assert np.abs(gp - wp).max() <= 1e-10 * max(1.0, np.abs(wp).max())   # the tolerance asked for


def test_unit_mass_kick_equals_the_reference_step(hgb):
    """mass None (= 1.0), at rest, dt = 1: v' = 0 + 1*a = a, which is what step() writes"""
    n = 3000
    pos = synth.uniform_cube(n, seed=12)
    with hgb.nbody.NBody(n) as nb:
        nb.upload(pos)
        nb.step(1)
        accel = nb.download()
        nb.upload(pos).set_dynamics(None, None, 1.0)
        nb.integrate(1)
        vel, newpos = nb.velocities(), nb.download()
    assert np.array_equal(vel, accel)
    assert np.array_equal(newpos, pos + accel)


def test_total_energy_vs_long_double(hgb):
    """K and U within 1e-12 relative of the long-double CPU sums, before and after 10 steps"""
    from oracle import port
    n = 6000
    pos, vel, mass = _case(n, seed=31)
    with hgb.nbody.NBody(n) as nb:
        nb.upload(pos).set_dynamics(vel, mass, 1e-3)
        k0, u0 = nb.total_energy()
        nb.integrate(10)
        k1, u1 = nb.total_energy()
        p1, v1 = nb.download(), nb.velocities()
    wk0, wu0 = port.nbody_total_energy(pos, vel, mass)
    wk1, wu1 = port.nbody_total_energy(p1, v1, mass)
    for got, want in ((k0, wk0), (u0, wu0), (k1, wk1), (u1, wu1)):
        assert abs(got - want) <= 1e-12 * abs(want), (got, want)
    e0, e1 = k0 + u0, k1 + u1
    assert abs(e1 - e0) <= 1e-3 * abs(e0)        # symplectic Euler, dt = 1e-3: energy stays put


def test_integrate_needs_dynamics_and_leaves_step_alone(hgb):
    n = 700
    pos, vel, mass = _case(n, seed=5)
    with hgb.nbody.NBody(n) as nb:
        nb.upload(pos)
        with pytest.raises(hgb.HgbError) as e:
            nb.integrate(1)
        assert e.value.code == hgb._abi.HGB_ERR_STATE
        with pytest.raises(hgb.HgbError):
            nb.velocities()
        nb.step(2)
        want = nb.download()
        nb.upload(pos).set_dynamics(vel, mass, 1e-3)
        nb.step(2)                               # the reference path ignores masses and velocities
        assert np.array_equal(nb.download(), want)
        assert np.array_equal(nb.velocities(), vel)
        assert nb.integrate(0) >= 0.0            # zero steps: nothing moves
        assert np.array_equal(nb.download(), want)


@pytest.mark.parametrize("exchange", ["nccl", "p2p"])
@pytest.mark.parametrize("n,steps", [(5000, 3), (70001, 2)])
def test_sharded_integrate_is_bit_identical(hgb, monkeypatch, exchange, n, steps):
    g = min(hgb.lib().hgb_device_count(), 8)
    if g < 2:
        if exchange == "nccl":
            pytest.skip("NCCL needs >= 2 devices (virtual shards use peer stores)")
        monkeypatch.setenv("HGB_VIRTUAL_GPUS", "4")
        g = 4
    kind = hgb.EXCHANGE_NCCL if exchange == "nccl" else hgb.EXCHANGE_P2P
    pos, vel, mass = _case(n, seed=n)
    with hgb.nbody.NBody(n, ngpus=1) as one:
        one.upload(pos).set_dynamics(vel, mass, 2e-3)
        one.integrate(steps)
        wp, wv, we = one.download(), one.velocities(), one.total_energy()
    for gg in sorted({2, 3 if g >= 3 else 2, g}):
        with hgb.nbody.NBody(n, ngpus=gg, exchange=kind) as nb:
            nb.upload(pos).set_dynamics(vel, mass, 2e-3)
            nb.integrate(steps)
            gp, gv, ge = nb.download(), nb.velocities(), nb.total_energy()
        assert np.array_equal(gp, wp) and np.array_equal(gv, wv), (exchange, gg)
        assert all(abs(a - b) <= 1e-12 * abs(b) for a, b in zip(ge, we))
