#!/usr/bin/env python
"""TEST INFRASTRUCTURE — generates the golden fixtures under tests/golden/ by running the *reference's own classes*
(oracle/_ref/ref_harness, built by oracle/build_ref.sh from /root/reference) and packing the stage dumps as .npz.

Runs only in the build container (needs /root/reference for the case data: chemistry/thermo/transport tables and the
shipped .cfg).  The fixtures it writes are committed; nothing at test time reads /root/reference.

It also prepares oracle/_ref/cases/<name>/ (git-ignored, travels to the GPU box) so that `bench.py --impl reference`
can run the compiled reference there.

usage: python oracle/make_golden.py [case ...]      (default: all)
       python oracle/make_golden.py prepare         (case directories only, fixtures untouched)
"""
import json
import os
import shutil
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = os.environ.get("SU2_REFERENCE", "/root/reference")
SHIPPED = os.path.join(REF, "Test_Cases", "TURBOLENT", "TURBOLENT_COMBUSTION")
CASES = os.path.join(HERE, "_ref", "cases")
HARNESS = os.path.join(HERE, "_ref", "ref_harness")
GOLDEN = os.path.join(ROOT, "tests", "golden")


def write_channel_su2(path, nx, ny, L=0.25, H=0.025, stretch=1.0):
    """Structured nx x ny point grid on [0,L]x[0,H]; every quad split into two triangles with alternating diagonal.
    Markers: inlet (x=0), outlet (x=L), lower_wall (y=0), upper_wall (y=H)."""
    xs = np.linspace(0.0, L, nx)
    t = np.linspace(0.0, 1.0, ny)
    if stretch != 1.0:  # symmetric wall clustering
        t = 0.5 * (1.0 + np.tanh(stretch * (2 * t - 1)) / np.tanh(stretch))
    ys = H * t
    idx = lambda i, j: j * nx + i
    with open(path, "w") as f:
        f.write("NDIME= 2\nNELEM= %d\n" % (2 * (nx - 1) * (ny - 1)))
        e = 0
        for j in range(ny - 1):
            for i in range(nx - 1):
                a, b, c, d = idx(i, j), idx(i + 1, j), idx(i + 1, j + 1), idx(i, j + 1)
                if (i + j) % 2 == 0:
                    tris = [(a, b, c), (a, c, d)]
                else:
                    tris = [(a, b, d), (b, c, d)]
                for tr in tris:
                    f.write("5 %d %d %d %d\n" % (tr[0], tr[1], tr[2], e)); e += 1
        f.write("NPOIN= %d\n" % (nx * ny))
        for j in range(ny):
            for i in range(nx):
                f.write("%.17g %.17g %d\n" % (xs[i], ys[j], idx(i, j)))
        f.write("NMARK= 4\n")
        f.write("MARKER_TAG= inlet\nMARKER_ELEMS= %d\n" % (ny - 1))
        for j in range(ny - 1): f.write("3 %d %d\n" % (idx(0, j), idx(0, j + 1)))
        f.write("MARKER_TAG= outlet\nMARKER_ELEMS= %d\n" % (ny - 1))
        for j in range(ny - 1): f.write("3 %d %d\n" % (idx(nx - 1, j), idx(nx - 1, j + 1)))
        f.write("MARKER_TAG= lower_wall\nMARKER_ELEMS= %d\n" % (nx - 1))
        for i in range(nx - 1): f.write("3 %d %d\n" % (idx(i, 0), idx(i + 1, 0)))
        f.write("MARKER_TAG= upper_wall\nMARKER_ELEMS= %d\n" % (nx - 1))
        for i in range(nx - 1): f.write("3 %d %d\n" % (idx(i, ny - 1), idx(i + 1, ny - 1)))


def derive_cfg(src, dst, repl):
    """Copy a shipped .cfg replacing `KEY= value` lines (keys matched case-sensitively before '=')."""
    done = set()
    out = []
    for line in open(src):
        s = line.strip()
        if s and not s.startswith("%") and "=" in s:
            key = s.split("=")[0].strip()
            if key in repl:
                if repl[key] is not None:
                    out.append("%s= %s\n" % (key, repl[key]))
                done.add(key)
                continue
        out.append(line)
    for k, v in repl.items():
        if k not in done and v is not None:
            out.append("%s= %s\n" % (k, v))
    open(dst, "w").writelines(out)


FLATPLATE = os.path.join(REF, "Test_Cases", "TURBOLENT", "TURBOLENT_FLAT_PLATE")


def prepare_case_dir(name, shipped_mesh=False, nx=0, ny=0, stretch=1.0, other_mixture=False):
    d = os.path.join(CASES, name)
    os.makedirs(d, exist_ok=True)
    for sub in ["Chemistry", "Mixture", "Thermo", "Transp"]:
        if not os.path.isdir(os.path.join(d, sub)):
            shutil.copytree(os.path.join(SHIPPED, sub), os.path.join(d, sub))
    for f in ["test_chem_first.txt", "test_chem_second.txt", "test_no_chem.txt"]:
        shutil.copy(os.path.join(SHIPPED, f), d)
    if other_mixture:   # the reference's other two mixtures (TURBOLENT_FLAT_PLATE): 3-species air (the shipped flat-plate cfg) and 5 species
        for sub in ["Mixture", "Thermo", "Transp"]:
            for f in os.listdir(os.path.join(FLATPLATE, sub)):
                shutil.copy(os.path.join(FLATPLATE, sub, f), os.path.join(d, sub, f))
        for f in ["test_air.txt", "test.txt"]:
            shutil.copy(os.path.join(FLATPLATE, f), d)
    if shipped_mesh:
        shutil.copy(os.path.join(SHIPPED, "mesh_stretched.su2"), d)
    else:
        write_channel_su2(os.path.join(d, "mesh.su2"), nx, ny, stretch=stretch)
    return d


CHANNEL_MARKERS = {
    "MARKER_ISOTHERMAL": "(upper_wall, 300.0, lower_wall, 600.0)",
    "MARKER_INLET": "( inlet, 300.0, 20.0, 1.0, 0.0, 0.0 )",
    "INLET_MASS_FRAC": "(inlet, 0.0, 0.0, 1.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0)",
    "MARKER_OUTLET": "( outlet, 101325.0)",
    "MESH_FILENAME": "mesh.su2",
    "MARKER_PLOTTING": "( upper_wall )", "MARKER_MONITORING": "( upper_wall )",
}

# harness environment per case (oracle/ref_harness.cpp: RX_PERTURB_TAMP, RX_POKE_ITER, RX_POKE_MODE)
ENVS = {
    "channel_lam_edge": {"RX_PERTURB_TAMP": "2000", "RX_POKE_ITER": "1", "RX_POKE_MODE": "abc"},
    "channel_lam_clip": {"RX_POKE_ITER": "1", "RX_POKE_MODE": "d"},
    "channel_rans_ign": {"RX_PERTURB_TAMP": "2000"},
    "channel_lam_inlet_total": {"RX_PERTURB_TAMP": "0", "RX_PERTURB_NOISE": "0"},   # total conditions need an interior state consistent with them
}

# name -> (case-dir kwargs, cfg replacements, nIter, dump_from, perturb_seed or None, lean fixture (drop intermediate residuals))
SPECS = {
    # config[0]: the jet-reactor case as shipped (REACTIVE_RANS + SST + PaSR, AUSM 1st order, WLS, explicit Euler)
    "jet_rans_shipped": (dict(shipped_mesh=True), {"RESTART_SOL": "NO", "EXT_ITER": "3"}, 3, 2, None, True),
    # same mesh and cfg, non-uniform synthetic state: exercises S-M diffusion, SST closure, PaSR, all BC kinds
    "jet_rans_perturbed": (dict(shipped_mesh=True), {"RESTART_SOL": "NO", "EXT_ITER": "2"}, 2, 1, 7, True),
    # small synthetic channel, laminar reactive NS + chemistry, explicit RK4-type, WLS gradients (config[1] in miniature)
    "channel_lam_rk": (dict(nx=41, ny=21, stretch=1.5),
                       dict(CHANNEL_MARKERS, RESTART_SOL="NO", EXT_ITER="2", KIND_TURB_MODEL="NONE", IGNITION="NO",
                            TIME_DISCRE_FLOW="RUNGE-KUTTA_EXPLICIT", RK_ALPHA_COEFF="( 0.66667, 0.66667, 1.000000 )",
                            CFL_NUMBER="0.2"), 2, 1, 11, False),
    # same channel, Green-Gauss gradients, explicit Euler
    "channel_lam_gg": (dict(nx=41, ny=21, stretch=1.5),
                       dict(CHANNEL_MARKERS, RESTART_SOL="NO", EXT_ITER="1", KIND_TURB_MODEL="NONE", IGNITION="NO",
                            NUM_METHOD_GRAD="GREEN_GAUSS"), 1, 0, 13, False),
    # same channel, RANS + PaSR
    "channel_rans": (dict(nx=41, ny=21, stretch=1.5),
                     dict(CHANNEL_MARKERS, RESTART_SOL="NO", EXT_ITER="2", IGNITION="NO"), 2, 1, 17, False),
    # same channel, laminar, the other two wall kinds of SURVEY.md 8(a10): heat-flux lower wall, Euler (slip) upper wall
    "channel_lam_walls": (dict(nx=41, ny=21, stretch=1.5),
                          dict(CHANNEL_MARKERS, MARKER_ISOTHERMAL="( NONE )", MARKER_HEATFLUX="( lower_wall, 5000.0 )",
                               MARKER_EULER="( upper_wall )", RESTART_SOL="NO", EXT_ITER="2", KIND_TURB_MODEL="NONE",
                               IGNITION="NO"), 2, 1, 19, False),
    # same channel, laminar, 2nd-order MUSCL reconstruction of [T,u,v,P] with the Venkatakrishnan limiter
    "channel_lam_muscl": (dict(nx=41, ny=21, stretch=1.5),
                          dict(CHANNEL_MARKERS, RESTART_SOL="NO", EXT_ITER="2", KIND_TURB_MODEL="NONE", IGNITION="NO",
                               SPATIAL_ORDER_FLOW="2ND_ORDER_LIMITER", SLOPE_LIMITER_FLOW="VENKATAKRISHNAN", CFL_NUMBER="0.2"),
                          2, 1, 23, False),
    # same with VISCOUS_LIMITER_FLOW= YES: the interior viscous flux forms its mean gradients of T, u, v from limited gradients (numerics_direct_reactive.cpp:1482-1505)
    "channel_lam_visclim": (dict(nx=41, ny=21, stretch=1.5),
                            dict(CHANNEL_MARKERS, RESTART_SOL="NO", EXT_ITER="2", KIND_TURB_MODEL="NONE", IGNITION="NO",
                                 SPATIAL_ORDER_FLOW="2ND_ORDER_LIMITER", SLOPE_LIMITER_FLOW="VENKATAKRISHNAN", VISCOUS_LIMITER_FLOW="YES", CFL_NUMBER="0.2"),
                            2, 1, 107, True),
    # same, BARTH_JESPERSEN branch of SetPrimitive_Limiter (as coded, order dependent)
    "channel_lam_bj": (dict(nx=41, ny=21, stretch=1.5),
                       dict(CHANNEL_MARKERS, RESTART_SOL="NO", EXT_ITER="2", KIND_TURB_MODEL="NONE", IGNITION="NO",
                            SPATIAL_ORDER_FLOW="2ND_ORDER_LIMITER", SLOPE_LIMITER_FLOW="BARTH_JESPERSEN", CFL_NUMBER="0.2"),
                       2, 1, 29, False),
    # inviscid multispecies reactive Euler solver (CReactiveEulerSolver itself): AUSM + source, slip walls, explicit RK
    "channel_euler": (dict(nx=41, ny=21, stretch=1.5),
                      dict(CHANNEL_MARKERS, PHYSICAL_PROBLEM="REACTIVE_EULER", MARKER_ISOTHERMAL="( NONE )",
                           MARKER_EULER="( upper_wall, lower_wall )", RESTART_SOL="NO", EXT_ITER="2", KIND_TURB_MODEL="NONE",
                           IGNITION="NO", TIME_DISCRE_FLOW="RUNGE-KUTTA_EXPLICIT", RK_ALPHA_COEFF="( 0.66667, 0.66667, 1.000000 )",
                           CFL_NUMBER="0.2"), 2, 1, 31, False),
    # RANS + PaSR with GREEN_GAUSS gradients: the flow path's GG branch under the SST closure, and CSolver::SetSolution_Gradient_GG
    # of (k, omega) for the turbulence postprocessing slice (tpost dump)
    "channel_rans_gg": (dict(nx=41, ny=21, stretch=1.5),
                        dict(CHANNEL_MARKERS, RESTART_SOL="NO", EXT_ITER="2", IGNITION="NO", NUM_METHOD_GRAD="GREEN_GAUSS"), 2, 1, 37, True),
    # edge cases of Cons2PrimVar no smooth state reaches (variable_direct_reactive.cpp:296-301, 656-688, 698-724): hot state
    # (T up to 2500 K), then at iteration 1 secant start values of 201 K and Tmax - 0.4 (exception -> bisection, secant not
    # converged -> bisection) and non-physical species densities (fallback to Solution_Old)
    "channel_lam_edge": (dict(nx=41, ny=21, stretch=1.5),
                         dict(CHANNEL_MARKERS, RESTART_SOL="NO", EXT_ITER="2", KIND_TURB_MODEL="NONE", IGNITION="NO",
                              TIME_DISCRE_FLOW="RUNGE-KUTTA_EXPLICIT", RK_ALPHA_COEFF="( 0.66667, 0.66667, 1.000000 )",
                              CFL_NUMBER="0.02"), 2, 1, 41, True),
    # CLIPPING_TEMPRATURE= YES (:728-739) with start values 20 % off at iteration 1
    "channel_lam_clip": (dict(nx=41, ny=21, stretch=1.5),
                         dict(CHANNEL_MARKERS, RESTART_SOL="NO", EXT_ITER="2", KIND_TURB_MODEL="NONE", IGNITION="NO",
                              CLIPPING_TEMPRATURE="YES", CFL_NUMBER="0.02"), 2, 1, 43, True),
    # RANS + PaSR on a burning state (T up to 2500 K) with the ignition overwrite switched off between iterations 1 and 2
    # (IGNITION_ITER= 2, solver_direct_reactive.cpp:1013-1024)
    "channel_rans_ign": (dict(nx=41, ny=21, stretch=1.5),
                         dict(CHANNEL_MARKERS, RESTART_SOL="NO", EXT_ITER="3", IGNITION="YES", IGNITION_ITER="2", CFL_NUMBER="0.02"), 3, 1, 47, True),
    # ---- the other two kinds of BC_Inlet (solver_direct_reactive.cpp:3283-3466): total conditions (the flat-plate cfg's kind) and mass flow
    "channel_lam_inlet_total": (dict(nx=41, ny=21, stretch=1.5),
                                dict(CHANNEL_MARKERS, INLET_TYPE="TOTAL_CONDITIONS", MARKER_INLET="( inlet, 500.5, 104000.0, 1.0, 0.0, 0.0 )",
                                     RESTART_SOL="NO", EXT_ITER="2", KIND_TURB_MODEL="NONE", IGNITION="NO",
                                     TIME_DISCRE_FLOW="RUNGE-KUTTA_EXPLICIT", RK_ALPHA_COEFF="( 0.66667, 0.66667, 1.000000 )", CFL_NUMBER="0.2"), 2, 1, 67, True),
    "channel_rans_inlet_massflow": (dict(nx=41, ny=21, stretch=1.5),
                                    dict(CHANNEL_MARKERS, INLET_TYPE="MASS_FLOW", MARKER_INLET="( inlet, 1.1, 25.0, 1.0, 0.0, 0.0 )",
                                         RESTART_SOL="NO", EXT_ITER="2", IGNITION="NO"), 2, 1, 71, True),
    # BC_Supersonic_Inlet / BC_Supersonic_Outlet (:2998-3214, :3681-3807) on the two ends of the channel, laminar
    "channel_lam_supersonic": (dict(nx=41, ny=21, stretch=1.5),
                               dict(CHANNEL_MARKERS, MARKER_INLET="( NONE )", MARKER_OUTLET="( NONE )",
                                    MARKER_SUPERSONIC_INLET="( inlet, 520.0, 101325.0, 700.0, 5.0, 0.0 )", MARKER_SUPERSONIC_OUTLET="( outlet )",
                                    RESTART_SOL="NO", EXT_ITER="2", KIND_TURB_MODEL="NONE", IGNITION="NO",
                                    TIME_DISCRE_FLOW="RUNGE-KUTTA_EXPLICIT", RK_ALPHA_COEFF="( 0.66667, 0.66667, 1.000000 )", CFL_NUMBER="0.05"), 2, 1, 73, True),
    # UNSTEADY_SIMULATION= TIME_STEPPING: every point advances with the global minimum time step (SetTime_Step :5262-5272); the Euler solver's
    # variant (:2121-2139) also swaps the CFL for UNST_CFL_NUMBER after the first call and takes UNST_TIMESTEP when that is zero
    "channel_lam_unsteady": (dict(nx=41, ny=21, stretch=1.5),
                             dict(CHANNEL_MARKERS, RESTART_SOL="NO", EXT_ITER="3", KIND_TURB_MODEL="NONE", IGNITION="NO", UNSTEADY_SIMULATION="TIME_STEPPING",
                                  UNST_CFL_NUMBER="0.1", UNST_TIMESTEP="1e-7", UNST_TIME="1.0",
                                  TIME_DISCRE_FLOW="RUNGE-KUTTA_EXPLICIT", RK_ALPHA_COEFF="( 0.66667, 0.66667, 1.000000 )", CFL_NUMBER="0.2"), 3, 0, 79, True),
    "channel_euler_unsteady": (dict(nx=41, ny=21, stretch=1.5),
                               dict(CHANNEL_MARKERS, PHYSICAL_PROBLEM="REACTIVE_EULER", MARKER_ISOTHERMAL="( NONE )", MARKER_EULER="( upper_wall, lower_wall )",
                                    RESTART_SOL="NO", EXT_ITER="3", KIND_TURB_MODEL="NONE", IGNITION="NO", UNSTEADY_SIMULATION="TIME_STEPPING",
                                    UNST_CFL_NUMBER="0.1", UNST_TIMESTEP="1e-7", UNST_TIME="1.0", CFL_NUMBER="0.2"), 3, 0, 83, True),
    "channel_euler_unsteady_fixed": (dict(nx=41, ny=21, stretch=1.5),
                                     dict(CHANNEL_MARKERS, PHYSICAL_PROBLEM="REACTIVE_EULER", MARKER_ISOTHERMAL="( NONE )", MARKER_EULER="( upper_wall, lower_wall )",
                                          RESTART_SOL="NO", EXT_ITER="2", KIND_TURB_MODEL="NONE", IGNITION="NO", UNSTEADY_SIMULATION="TIME_STEPPING",
                                          UNST_CFL_NUMBER="0.0", UNST_TIMESTEP="2e-8", UNST_TIME="1.0", CFL_NUMBER="0.2"), 2, 0, 89, True),
    # ---- dual time stepping, explicit pseudo-time iterations with a fixed physical time step: 2nd order on the laminar RK case,
    # 1st order on the Euler case with the physical step taken from UNST_CFL_NUMBER on iteration 0; the time levels are shifted after every second pseudo-time iteration (RX_DUAL_INNER)
    "channel_lam_dualtime2": (dict(nx=41, ny=21, stretch=1.5),
                              dict(CHANNEL_MARKERS, RESTART_SOL="NO", EXT_ITER="3", KIND_TURB_MODEL="NONE", IGNITION="NO",
                                   UNSTEADY_SIMULATION="DUAL_TIME_STEPPING-2ND_ORDER", UNST_CFL_NUMBER="0.0", UNST_TIMESTEP="5e-8", UNST_TIME="1.0",
                                   TIME_DISCRE_FLOW="RUNGE-KUTTA_EXPLICIT", RK_ALPHA_COEFF="( 0.66667, 0.66667, 1.000000 )", CFL_NUMBER="0.2"), 3, 1, 97, True),
    "channel_euler_dualtime1": (dict(nx=41, ny=21, stretch=1.5),
                                dict(CHANNEL_MARKERS, PHYSICAL_PROBLEM="REACTIVE_EULER", MARKER_ISOTHERMAL="( NONE )", MARKER_EULER="( upper_wall, lower_wall )",
                                     RESTART_SOL="NO", EXT_ITER="4", KIND_TURB_MODEL="NONE", IGNITION="NO",
                                     UNSTEADY_SIMULATION="DUAL_TIME_STEPPING-1ST_ORDER", UNST_CFL_NUMBER="1.5", UNST_TIMESTEP="4e-8", UNST_TIME="1.0",
                                     CFL_NUMBER="0.2"), 4, 0, 101, True),
    # RANS: the flow path of a dual-time run with the reference's own (host) SST solver, whose SetResidual_DualTime reads the flow
    # solver's density at the three time levels (solver_direct_turbulent.cpp:778-780)
    "channel_rans_dualtime2": (dict(nx=41, ny=21, stretch=1.5),
                               dict(CHANNEL_MARKERS, RESTART_SOL="NO", EXT_ITER="4", IGNITION="NO",
                                    UNSTEADY_SIMULATION="DUAL_TIME_STEPPING-2ND_ORDER", UNST_CFL_NUMBER="0.0", UNST_TIMESTEP="5e-8", UNST_TIME="1.0"), 4, 2, 103, True),
    # ---- species-count generality: the reference's other mixtures (Test_Cases/TURBOLENT/TURBOLENT_FLAT_PLATE) on the same channel
    # 3-species air (O2, CO2, N2; the mixture of the shipped flat-plate cfg, no reactions), laminar, explicit RK
    "channel_air3_lam": (dict(nx=41, ny=21, stretch=1.5, other_mixture=True),
                         dict(CHANNEL_MARKERS, CONFIG_LIB_FILE="test_air.txt", FREESTREAM_MASS_FRAC="(0.2197, 0.0302, 0.7501)",
                              SPECIES_ORDER="(O2, CO2, N2)", INLET_MASS_FRAC="(inlet, 0.2197, 0.0302, 0.7501)", FUEL_INDEX="1", OXIDIZER_INDEX="0",
                              RESTART_SOL="NO", EXT_ITER="2", KIND_TURB_MODEL="NONE", IGNITION="NO",
                              TIME_DISCRE_FLOW="RUNGE-KUTTA_EXPLICIT", RK_ALPHA_COEFF="( 0.66667, 0.66667, 1.000000 )", CFL_NUMBER="0.2"), 2, 1, 53, False),
    # the same mixture, RANS (SST closure with a 3 x 3 molar->mass operator)
    "channel_air3_rans": (dict(nx=41, ny=21, stretch=1.5, other_mixture=True),
                          dict(CHANNEL_MARKERS, CONFIG_LIB_FILE="test_air.txt", FREESTREAM_MASS_FRAC="(0.2197, 0.0302, 0.7501)",
                               SPECIES_ORDER="(O2, CO2, N2)", INLET_MASS_FRAC="(inlet, 0.2197, 0.0302, 0.7501)", FUEL_INDEX="1", OXIDIZER_INDEX="0",
                               RESTART_SOL="NO", EXT_ITER="2", IGNITION="NO"), 2, 1, 59, False),
    # 5 species (C4H6, H2O, O2, CO, CO2; test.txt), laminar, MUSCL + Venkatakrishnan
    "channel_mix5_lam": (dict(nx=41, ny=21, stretch=1.5, other_mixture=True),
                         dict(CHANNEL_MARKERS, CONFIG_LIB_FILE="test.txt", FREESTREAM_MASS_FRAC="(0.0, 0.0, 1.0, 0.0, 0.0)",
                              SPECIES_ORDER="(C4H6, H2O, O2, CO, CO2)", INLET_MASS_FRAC="(inlet, 0.0, 0.0, 1.0, 0.0, 0.0)",
                              RESTART_SOL="NO", EXT_ITER="2", KIND_TURB_MODEL="NONE", IGNITION="NO",
                              SPATIAL_ORDER_FLOW="2ND_ORDER_LIMITER", SLOPE_LIMITER_FLOW="VENKATAKRISHNAN", CFL_NUMBER="0.2"), 2, 1, 61, False),
}


def prepare_case(name):
    """Case directory + derived cfg only (what tests/test_gpu_dropin.py and bench.py need on the GPU box)."""
    kw, repl = SPECS[name][0], SPECS[name][1]
    d = prepare_case_dir(name, **kw)
    derive_cfg(os.path.join(SHIPPED, "my_combustion_first_chem_PaSR.cfg"), os.path.join(d, "case.cfg"), repl)
    return d


def run_case(name):
    kw, repl, nit, dump_from, seed, lean = SPECS[name]
    d = prepare_case(name)
    out = os.path.join(d, "dump")
    shutil.rmtree(out, ignore_errors=True)
    cmd = [HARNESS, "case.cfg", str(nit), out, str(dump_from)] + ([str(seed)] if seed is not None else [])
    log = subprocess.run(cmd, cwd=d, env=dict(os.environ, **ENVS.get(name, {})), stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    tail = "\n".join([l for l in log.stdout.splitlines() if "poke" in l or "bisection" in l][:6] + log.stdout.splitlines()[-6:])
    if log.returncode != 0:
        raise RuntimeError("harness failed for %s:\n%s" % (name, tail))
    print(tail)
    arrs = {}
    for f in sorted(os.listdir(out)):
        if f.endswith(".npy"):
            k = f[:-4]
            # the "prev_" dumps are only needed for T (secant start value); aux/grad copies are dropped to keep fixtures small
            if k.endswith("prev_aux") or k.endswith("prev_grad") or k.endswith("out_grad"):
                continue
            if lean and k.split("_", 2)[-1] in ("res_conv", "res_visc", "res_src"):
                continue
            a = np.load(os.path.join(out, f))
            if k.endswith("out_V") and os.path.exists(os.path.join(out, k[:-5] + "tpost.npy")):
                # flow fields CTurbSSTSolver::Postprocessing reads (rho, laminar viscosity, strain magnitude), kept even for lean fixtures
                aux = np.load(os.path.join(out, k[:-1] + "aux.npy"))
                arrs[k[:-5] + "tpost_flow"] = np.ascontiguousarray(np.stack([a[:, 4], aux[:, 2], aux[:, 5]], axis=1))
            if lean and k.split("_", 2)[-1] in ("out_V", "out_aux"):
                continue
            if k.endswith("prev_V"):
                a = np.ascontiguousarray(a[:, :1])
            arrs[k] = a
    meta = open(os.path.join(out, "meta.json")).read()
    json.loads(meta)
    arrs["meta_json"] = np.frombuffer(meta.encode(), dtype=np.uint8)
    os.makedirs(GOLDEN, exist_ok=True)
    path = os.path.join(GOLDEN, name + ".npz")
    np.savez_compressed(path, **arrs)
    print("wrote %s  (%.2f MB)" % (path, os.path.getsize(path) / 1e6))
    shutil.rmtree(out, ignore_errors=True)


# long runs for the converged-field / residual-history parity (north_star: fields and history within 1e-10): same case directories
# and cfgs as the stage fixtures of the same name, more iterations, checkpoints instead of stage dumps.
# name -> (stage-fixture name, cfg overrides, harness environment, iterations, perturb seed or None, checkpoints)
# The laminar channel: 3-stage RK at CFL 0.02 from a cool perturbed state (temperature bump of 300 K) — at the stage fixture's CFL 0.2
# the reference itself diverges within five iterations, and a hot state ignites; this one relaxes steadily.
LONG = {
    "long_channel_lam_rk": ("channel_lam_rk", {"CFL_NUMBER": "0.02"}, {"RX_PERTURB_TAMP": "300"}, 2000, 11, [0, 1, 10, 100, 500, 1000, 2000]),
    "long_jet_rans_shipped": ("jet_rans_shipped", {}, {}, 500, None, [0, 1, 10, 100, 250, 500]),
}


def run_long(name):
    base, repl, envx, nit, seed, ckpt = LONG[name]
    d = prepare_case(base)
    cfg = "case_long.cfg"
    derive_cfg(os.path.join(d, "case.cfg"), os.path.join(d, cfg), repl)
    out = os.path.join(d, "dump_long")
    shutil.rmtree(out, ignore_errors=True)
    cmd = [HARNESS, cfg, str(nit), out, str(10 ** 9)] + ([str(seed)] if seed is not None else [])
    env = dict(os.environ, RX_CKPT=",".join(map(str, ckpt)), RX_QUIET="1", **envx)
    log = subprocess.run(cmd, cwd=d, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if log.returncode != 0:
        raise RuntimeError("harness failed for %s:\n%s" % (name, log.stdout[-2000:]))
    arrs = {f[:-4]: np.load(os.path.join(out, f)) for f in sorted(os.listdir(out)) if f.startswith(("ckpt", "history_rms"))}
    arrs["meta_json"] = np.frombuffer(json.dumps(dict(base=base, cfg=repl, env=envx, nIter=nit, seed=seed, ckpt=ckpt)).encode(), dtype=np.uint8)
    path = os.path.join(GOLDEN, name + ".npz")
    np.savez_compressed(path, **arrs)
    print("wrote %s  (%.2f MB)" % (path, os.path.getsize(path) / 1e6))
    shutil.rmtree(out, ignore_errors=True)


# restart files written by the reference's own COutput (MergeCoordinates / MergeSolution / SetRestart through the harness) after the
# last iteration of the stage fixture of the same name: the golden bytes for rx_restart_write
RESTART = ["channel_lam_rk", "channel_rans"]


def run_restart_files():
    arrs = {}
    for name in RESTART:
        kw, repl, nit, dump_from, seed, lean = SPECS[name]
        d = prepare_case(name)
        out = os.path.join(d, "dump_rst")
        shutil.rmtree(out, ignore_errors=True)
        cmd = [HARNESS, "case.cfg", str(nit), out, str(10 ** 9)] + ([str(seed)] if seed is not None else [])
        log = subprocess.run(cmd, cwd=d, env=dict(os.environ, RX_RESTART_OUT="1", RX_TECPLOT_OUT="1", **ENVS.get(name, {})), stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if log.returncode != 0 or "HARNESS wrote" not in log.stdout:
            raise RuntimeError("harness failed for %s:\n%s" % (name, log.stdout[-2000:]))
        arrs[name] = np.frombuffer(open(os.path.join(d, "restart_flow.dat"), "rb").read(), dtype=np.uint8)
        os.remove(os.path.join(d, "restart_flow.dat"))
        # the Tecplot ASCII volume file of the same state (COutput::SetTecplotASCII): golden bytes for rx_tecplot_write
        arrs[name + "_tecplot"] = np.frombuffer(open(os.path.join(d, "flow.dat"), "rb").read(), dtype=np.uint8)
        os.remove(os.path.join(d, "flow.dat"))
        shutil.rmtree(out, ignore_errors=True)
    path = os.path.join(GOLDEN, "restart_files.npz")
    np.savez_compressed(path, **arrs)
    print("wrote %s  (%.2f MB)" % (path, os.path.getsize(path) / 1e6))


def prepare_bench_case(nx=151, ny=151):
    """Case directory for `bench.py --impl reference` / cpu_baseline: the bench workload (laminar reactive NS, 3-stage RK,
    WLS gradients, AUSM 1st order) at a size the serial CPU reference finishes in seconds per iteration."""
    d = prepare_case_dir("bench_ref", nx=nx, ny=ny, stretch=1.0)
    derive_cfg(os.path.join(SHIPPED, "my_combustion_first_chem_PaSR.cfg"), os.path.join(d, "case.cfg"),
               dict(CHANNEL_MARKERS, RESTART_SOL="NO", EXT_ITER="100000", KIND_TURB_MODEL="NONE", IGNITION="NO",
                    TIME_DISCRE_FLOW="RUNGE-KUTTA_EXPLICIT", RK_ALPHA_COEFF="( 0.66667, 0.66667, 1.000000 )", CFL_NUMBER="0.1"))
    print("prepared %s" % d)


def prepare_dropin_case(nx=501, ny=501):
    """Case directory for `bench.py --workload dropin`: the bench cfg on a 251 001-point channel — the real cost of the C++ drop-in
    (ref_harness_gpu at SU2_RX_MIRROR=0 / 1) next to the CPU reference on the same mesh."""
    d = prepare_case_dir("dropin_250k", nx=nx, ny=ny, stretch=1.0)
    derive_cfg(os.path.join(SHIPPED, "my_combustion_first_chem_PaSR.cfg"), os.path.join(d, "case.cfg"),
               dict(CHANNEL_MARKERS, RESTART_SOL="NO", EXT_ITER="100000", KIND_TURB_MODEL="NONE", IGNITION="NO",
                    TIME_DISCRE_FLOW="RUNGE-KUTTA_EXPLICIT", RK_ALPHA_COEFF="( 0.66667, 0.66667, 1.000000 )", CFL_NUMBER="0.1"))
    print("prepared %s" % d)


if __name__ == "__main__":
    if sys.argv[1:] == ["bench_ref"]:
        prepare_bench_case()
        sys.exit(0)
    if sys.argv[1:] == ["prepare"]:   # case directories only; the committed fixtures are left alone
        for n in SPECS:
            prepare_case(n)
        for n in RESTART:   # cfg that starts from restart_flow.dat (round trip of rx_restart_write through the reference's readers)
            derive_cfg(os.path.join(CASES, n, "case.cfg"), os.path.join(CASES, n, "case_restart.cfg"), {"RESTART_SOL": "YES"})
        for base, repl, _e, _n, _s, _c in LONG.values():
            derive_cfg(os.path.join(CASES, base, "case.cfg"), os.path.join(CASES, base, "case_long.cfg"), repl)
        prepare_bench_case()
        prepare_dropin_case()
        sys.exit(0)
    names = sys.argv[1:] or list(SPECS)
    for n in names:
        print("== %s" % n)
        if n == "restart_files":
            run_restart_files()
        elif n in LONG:
            run_long(n)
        else:
            run_case(n)
    prepare_bench_case()
