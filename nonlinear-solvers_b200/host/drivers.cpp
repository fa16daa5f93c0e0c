// drivers.cpp -- the reference's *_driver command lines (SURVEY.md 8b "B1") on top of the
// B200 engine.  One table-driven program; the executable's basename (or `nls_driver <name>`)
// selects the driver, so the Python launchers' `--exe` keeps working unchanged.
//
// Kept from the reference, per driver (file:line in the table below):
//   positional argv and usage-on-wrong-argc (exit 1), dx = 2*Lx/(nx-1), dt = T/nt,
//   freq = nt/num_snapshots, loop i = 1..nt-1, boundary copy after every step, snapshot rule
//   i % freq == 0 && i/freq < num_snapshots, snapshot 0 = initial data, .npy shapes
//   (ny,nx) / (nz,ny,nx) in and (ns,[nz,]ny,nx) out, default m when no file is given,
//   the sign of the operator (+L or -L) and the hard-coded Krylov dimension.
// Optional trailing flags (not in the reference; ignored by its launchers):
//   --m K   Krylov dimension      --device D   first CUDA device ordinal
//   --diag d.npy  also write [mass, energy] of every stored snapshot, shape (num_snapshots, 2)
//                 [(B, num_snapshots, 2)]: NLS_RED_MASS / NLS_RED_ENERGY, the quantities of
//                 compare_utils_complex_2d.py:134-151 and energy_sge.py:6-14, reduced on the device
//   --downsample N  store every snapshot resampled on the device to N x N [x N] points with the launcher's
//                 downsample_interpolation (finalized_scripts/downsampling.py:68-104, :215-258); the
//                 trajectory files then have shape ([B,] num_snapshots, N, N[, N])
//   --downsample-fft N  the same with the launcher's downsample_fft (downsampling.py:10-34; 2D drivers): spectral crop
//   --devices G   spread an ensemble file over devices D..D+G-1 (one host thread each)
// Ensemble input: when u0.npy [and v0.npy] carry one extra leading dimension
// `(B, [nz,] ny, nx)`, all B trajectories advance in batched handles and the trajectory files
// get the shape `(B, num_snapshots, ...)`; `nlse_ensemble_driver` is the cubic-NLSE alias for it.
#include "../../include/nls/solvers.hpp"
#include "npy.hpp"

#include <chrono>
#include <cstdlib>
#include <iostream>
#include <libgen.h>
#include <thread>
#include <algorithm>

using nls_host::read_npy;
using nls_host::write_npy;
using cplx = std::complex<double>;

namespace {

struct Spec {
  const char *name;
  int dim;          // 2 or 3
  int equation;     // NLS_EQ_*
  int integrator;   // NLS_INT_*
  int nscal;        // extra scalars after the box lengths: cq -> s1 s2, saturating -> kappa
  bool fields_req;  // true: "... m.npy c.npy" both required (gen-2 device drivers, all 3D drivers)
  double m_default; // value of m(x) when the optional m file is absent
  double op_sign;   // +1: L, -1: -L
  uint32_t krylov;  // Krylov dimension hard-coded by that driver
  const char *ref;  // reference source
  uint32_t flags = 0; // NLS_FLAG_*
};

const Spec kSpecs[] = {
    // host (CPU) drivers, nlsolvers/host/drivers/
    {"nlse_cubic", 2, NLS_EQ_NLSE_CUBIC, NLS_INT_SS2, 0, false, 0.0, 1, 10, "nlse_cubic_driver.cpp:18-117"},
    {"nlse_cubic_quintic", 2, NLS_EQ_NLSE_CQ, NLS_INT_SS2, 2, false, 0.0, 1, 10, "nlse_cubic_quintic_driver.cpp:18-118"},
    {"nlse_saturating", 2, NLS_EQ_NLSE_SAT, NLS_INT_SS2, 1, false, 0.0, 1, 10, "nlse_saturating_driver.cpp:18-117"},
    {"nlse_cubic_host_3d", 3, NLS_EQ_NLSE_CUBIC, NLS_INT_SS2, 0, true, 0.0, 1, 10, "nlse_cubic_driver_3d.cpp:21-140"},
    {"sg_single", 2, NLS_EQ_SGE, NLS_INT_GAUTSCHI, 0, false, 1.0, 1, 10, "sg_single_driver.cpp:20-126"},
    {"sg_single_sv", 2, NLS_EQ_SGE, NLS_INT_SV, 0, false, 1.0, 1, 10, "sg_single_driver_sv.cpp:20-126"},
    {"sg_double", 2, NLS_EQ_SGE_DOUBLE, NLS_INT_GAUTSCHI, 0, false, 1.0, -1, 10, "sg_double_driver.cpp:20-125"},
    {"sg_hyperbolic", 2, NLS_EQ_SGE_SINH, NLS_INT_GAUTSCHI, 0, false, 1.0, -1, 10, "sg_hyperbolic_driver.cpp:20-125"},
    {"kg", 2, NLS_EQ_KGE, NLS_INT_GAUTSCHI, 0, false, 1.0, -1, 10, "kg_driver.cpp:20-126"},
    {"kg_sv", 2, NLS_EQ_KGE, NLS_INT_SV, 0, false, 1.0, 1, 10, "kg_sv_driver.cpp:20-126"},
    {"phi4", 2, NLS_EQ_PHI4, NLS_INT_GAUTSCHI, 0, false, 1.0, -1, 10, "phi4_driver.cpp:20-123"},
    {"kg_driver_3d", 3, NLS_EQ_KGE, NLS_INT_GAUTSCHI, 0, true, 1.0, -1, 10, "kg_driver_3d.cpp:20-146"},
    {"kg_sv_driver_3d", 3, NLS_EQ_KGE, NLS_INT_SV, 0, true, 1.0, 1, 10, "kg_sv_driver_3d.cpp:20-146"},
    // in-process ensemble alias (N3): the nlse_cubic contract with m == 1 when no m file is given,
    // like the trajectories of device/nlse_driver_omp.cpp:103-121 (u_t = i(Lu + |u|^2 u))
    {"nlse_ensemble_driver", 2, NLS_EQ_NLSE_CUBIC, NLS_INT_SS2, 0, false, 1.0, 1, 10, "nlse_driver_omp.cpp:56-121"},
    // sEWI ("Gautschi" NLSE): one SS2 start-up step, then the two-step integrator
    {"nlse_cubic_gautschi", 2, NLS_EQ_NLSE_CUBIC, NLS_INT_SEWI, 0, false, 0.0, 1, 10, "nlse_cubic_gautschi_driver.cpp:19-150",
     NLS_FLAG_SEWI_HOST2D},
    {"nlse_cubic_gautschi_3d", 3, NLS_EQ_NLSE_CUBIC, NLS_INT_SEWI, 0, true, 0.0, 1, 10, "nlse_cubic_gautschi_driver_3d.cpp:22-150"},
    {"nlse_sewi_2d_dev", 2, NLS_EQ_NLSE_CUBIC, NLS_INT_SEWI, 0, true, 0.0, 1, 25, "nlse_cubic_sewi_driver_2d.cpp:19-125"},
    {"nlse_sewi_3d_dev", 3, NLS_EQ_NLSE_CUBIC, NLS_INT_SEWI, 0, true, 0.0, 1, 15, "nlse_cubic_sewi_driver_3d.cpp:19-133"},
    // gen-2 device drivers, nlsolvers/device/drivers/ (names from its CMakeLists.txt:55-73)
    {"nlse_2d_dev", 2, NLS_EQ_NLSE_CUBIC, NLS_INT_SS2, 0, true, 0.0, 1, 20, "nlse_cubic_driver_2d.cpp:19-125"},
    {"nlse_3d_dev", 3, NLS_EQ_NLSE_CUBIC, NLS_INT_SS2, 0, true, 0.0, 1, 25, "nlse_cubic_driver_3d.cpp:19-133"},
    {"nlse_cubic_quintic_dev", 2, NLS_EQ_NLSE_CQ, NLS_INT_SS2, 2, false, 1.0, 1, 15, "nlse_cubic_quintic_driver_dev.cpp:16-118"},
    {"kg_gautschi_2d_dev", 2, NLS_EQ_KGE, NLS_INT_GAUTSCHI, 0, true, 1.0, -1, 10, "kg_driver_dev_2d.cpp:20-165"},
    {"kg_sv_2d_dev", 2, NLS_EQ_KGE, NLS_INT_SV, 0, true, 1.0, 1, 10, "kg_sv_driver_dev_2d.cpp:20-162"},
    {"kg_gautschi_3d_dev", 3, NLS_EQ_KGE, NLS_INT_GAUTSCHI, 0, true, 1.0, -1, 10, "kg_driver_dev_3d.cpp:19-170"},
    {"kg_sv_3d_dev", 3, NLS_EQ_KGE, NLS_INT_SV, 0, true, 1.0, 1, 10, "kg_sv_driver_dev_3d.cpp"},
    {"sge_sv_2d_dev", 2, NLS_EQ_SGE, NLS_INT_SV, 0, true, 1.0, 1, 10, "sg_sv_driver_dev_2d.cpp:20-162"},
    {"sge_sv_3d_dev", 3, NLS_EQ_SGE, NLS_INT_SV, 0, true, 1.0, 1, 10, "sg_sv_driver_dev_3d.cpp"},
};

bool is_nlse(const Spec &s) { return s.equation <= NLS_EQ_NLSE_SAT; }

int usage(const Spec &s, const char *argv0) {
  std::cerr << "Usage: " << argv0 << (s.dim == 3 ? " nx ny nz Lx Ly Lz" : " nx ny Lx Ly");
  if (s.nscal == 2) std::cerr << " sigma1 sigma2";
  if (s.nscal == 1) std::cerr << " kappa";
  if (is_nlse(s)) std::cerr << " input_u0.npy output_traj.npy";
  else std::cerr << " input_u0.npy input_v0.npy output_traj.npy output_vel.npy";
  std::cerr << " T nt num_snapshots" << (s.fields_req ? " input_m.npy input_c.npy" : " [input_m.npy]")
            << " [--m K] [--device D] [--devices G] [--diag d.npy] [--downsample N | --downsample-fft N]\n"
            << "  (u0/v0 may be ensemble files of shape (B, ...): output is (B, num_snapshots, ...))\n";
  return 1;
}

template <class T>
bool load_checked(const std::string &file, const std::vector<uint64_t> &want, std::vector<T> &out, const char *what) {
  std::vector<uint64_t> shape;
  out = read_npy<T>(file, shape);
  if (shape == want) return true;
  std::cerr << "Error: " << what << " array dimensions mismatch\nExpected:";
  for (auto v : want) std::cerr << " " << v;
  std::cerr << "\nGot:";
  for (auto v : shape) std::cerr << " " << v;
  std::cerr << "\n";
  return false;
}

// Shape check that also accepts a leading ensemble dimension: `(B, [nz,] ny, nx)` holds B
// independent initial conditions (the reference runs one process, or one OpenMP
// thread with its own solver, per trajectory: device/submit_nlse.py:117-129,
// device/nlse_driver_omp.cpp:103-121).  B = 0 on return means "not an ensemble file".
template <class T>
bool load_state(const std::string &file, const std::vector<uint64_t> &want, std::vector<T> &out, const char *what,
                uint64_t &B) {
  std::vector<uint64_t> shape;
  out = read_npy<T>(file, shape);
  if (shape == want) { B = 0; return true; }
  if (shape.size() == want.size() + 1 && shape[0] >= 1 && std::equal(want.begin(), want.end(), shape.begin() + 1)) {
    B = shape[0];
    return true;
  }
  std::cerr << "Error: " << what << " array dimensions mismatch\nExpected:";
  for (auto v : want) std::cerr << " " << v;
  std::cerr << "\nGot:";
  for (auto v : shape) std::cerr << " " << v;
  std::cerr << "\n";
  return false;
}

// The time loop of every driver (nlse_cubic_driver.cpp:96-105, sg_single_driver.cpp:104-121) for
// one trajectory or for an ensemble file.  The B trajectories are block-partitioned over the
// devices (`[g*B/G, (g+1)*B/G)`), one host thread and one batched handle per
// device, no communication.  Output: `(ns, ...)` for a plain input, `(B, ns, ...)` for an
// ensemble input, trajectory b identical to what the single-trajectory driver writes for IC b.
template <class T>
bool run_ensemble(const Spec &s, const nls::Grid &g, uint32_t krylov, double dt, const double *sc, uint32_t nt,
                  uint32_t ns, uint32_t freq, const std::vector<uint64_t> &want, const std::vector<double> &m,
                  const std::string &in_u, const std::string &in_v, const std::string &out_u,
                  const std::string &out_v, const std::vector<int> &devices, uint64_t &B_out,
                  const std::string &diag_file, uint32_t ds, bool ds_fft) {
  const bool real = !in_v.empty();
  const size_t n = g.n();
  std::vector<T> u0, v0;
  uint64_t Bu = 0, Bv = 0;
  if (!load_state(in_u, want, u0, "Input", Bu)) return false;
  if (real && !load_state(in_v, want, v0, "Input velocity", Bv)) return false;
  if (real && Bu != Bv) {
    std::cerr << "Error: u0 and v0 hold a different number of trajectories (" << Bu << " vs " << Bv << ")\n";
    return false;
  }
  const uint64_t B = Bu ? Bu : 1;
  B_out = B;
  if (devices.size() > 1) { // a single device is left to nls_create to validate, like every other driver
    const int have = std::max(0, nls_device_count());
    if (devices.back() >= have) {
      std::cerr << "Error: --device " << devices.front() << " --devices " << devices.size() << " but " << have
                << " CUDA device(s) are visible\n";
      return false;
    }
  }
  // --downsample N: every stored snapshot (snapshot 0 included) is resampled on the device to N^dim
  // points (nls_get_field_resampled = downsampling.py:68-104) and only that crosses PCIe / hits the file
  const uint32_t dtz = g.nz > 1 ? ds : 1;
  const size_t no = ds ? (size_t)ds * ds * dtz : n; // scalars per stored snapshot
  std::vector<uint64_t> oshape = want;
  if (ds) oshape.assign(want.size(), ds);
  oshape.insert(oshape.begin(), ns);
  if (Bu) oshape.insert(oshape.begin(), B);
  // the trajectory file images live in page-locked memory: every snapshot leaves the device through the
  // asynchronous egress ring (nls_get_field[_resampled]_async) straight into its (b, snapshot) slot while the next
  // segment's steps run (the reference blocks on a cudaMemcpy per snapshot, nlse_dev.hpp:323-334)
  nls::Pinned<T> tu((size_t)B * ns * no), tv(real ? (size_t)B * ns * no : 0);
  if (!ds)
    for (uint64_t b = 0; b < B; ++b) {
      std::copy(u0.begin() + b * n, u0.begin() + (b + 1) * n, tu.data() + b * ns * n);
      if (real) std::copy(v0.begin() + b * n, v0.begin() + (b + 1) * n, tv.data() + b * ns * n);
    }

  // --diag: [mass, energy] of every stored snapshot, reduced on the device (nls_reduce)
  const bool diag = !diag_file.empty();
  std::vector<double> dg(diag ? (size_t)B * ns * 2 : 0);
  const size_t G = std::min<size_t>(devices.size(), B);
  std::vector<std::string> errors(G);
  auto worker = [&](size_t gi) {
    try {
      const uint64_t b0 = gi * B / G, b1 = (gi + 1) * B / G, nb = b1 - b0;
      nls::Handle h(g, s.equation, s.integrator, krylov, dt, sc[0], sc[1], sc[0], (uint32_t)nb, devices[gi], s.flags);
      h.set(NLS_FIELD_M, m.data(), n * 8);
      h.set(NLS_FIELD_U, u0.data() + b0 * n, nb * n * sizeof(T));
      if (real) h.set(NLS_FIELD_V, v0.data() + b0 * n, nb * n * sizeof(T));
      std::vector<double> red(nb);
      auto diagnose = [&](size_t snap) {
        if (!diag) return;
        for (int q = 0; q < 2; ++q) {
          red = h.reduce(q == 0 ? NLS_RED_MASS : NLS_RED_ENERGY);
          for (uint64_t b = 0; b < nb; ++b) dg[((b0 + b) * ns + snap) * 2 + q] = red[b];
        }
      };
      const size_t tstride = (size_t)ns * no * sizeof(T); // bytes between consecutive trajectories in the file image
      auto store = [&](size_t snap) {
        diagnose(snap);
        T *du = tu.data() + (b0 * ns + snap) * no;
        T *dv = real ? tv.data() + (b0 * ns + snap) * no : nullptr;
        if (ds && ds_fft) {
          h.get_fft_resampled_async(NLS_FIELD_U, ds, ds, du, nb * no * sizeof(T), tstride);
          if (real) h.get_fft_resampled_async(NLS_FIELD_V, ds, ds, dv, nb * no * sizeof(T), tstride);
        } else if (ds) {
          h.get_resampled_async(NLS_FIELD_U, ds, ds, dtz, du, nb * no * sizeof(T), tstride);
          if (real) h.get_resampled_async(NLS_FIELD_V, ds, ds, dtz, dv, nb * no * sizeof(T), tstride);
        } else {
          h.get_async(NLS_FIELD_U, du, nb * n * sizeof(T), tstride);
          if (real) h.get_async(NLS_FIELD_V, dv, nb * n * sizeof(T), tstride);
        }
      };
      if (ds) store(0);
      else diagnose(0);
      // run up to the next snapshot step in one asynchronous call; same steps, same snapshots
      for (uint32_t i = 1; i < nt;) {
        uint32_t next = (i / freq + (i % freq ? 1 : 0)) * freq; // first multiple of freq >= i
        if (next >= nt) next = nt - 1;
        h.step(dt, next - i + 1);
        i = next + 1;
        if (next % freq == 0 && next / freq < ns) store(next / freq);
      }
      h.wait(); // steps and snapshot transfers
    } catch (const std::exception &e) { errors[gi] = e.what(); }
  };
  if (G == 1) worker(0);
  else {
    std::vector<std::thread> th;
    for (size_t gi = 0; gi < G; ++gi) th.emplace_back(worker, gi);
    for (auto &t : th) t.join();
  }
  for (const auto &e : errors)
    if (!e.empty()) throw std::runtime_error(e); // CUDA/engine errors abort like the reference
  write_npy(out_u, tu.data(), oshape);
  if (real) write_npy(out_v, tv.data(), oshape);
  if (diag) {
    std::vector<uint64_t> dshape = {ns, 2};
    if (Bu) dshape.insert(dshape.begin(), B);
    write_npy(diag_file, dg.data(), dshape);
  }
  return true;
}

int run(const Spec &s, int argc, char **argv) {
  // strip the optional trailing flags
  uint32_t krylov = s.krylov;
  int device = 0, ndev = 1;
  std::string diag_file;
  uint32_t ds = 0;
  bool ds_fft = false;
  if (const char *e = std::getenv("NLS_KRYLOV_M")) krylov = (uint32_t)std::stoul(e);
  std::vector<char *> av(argv, argv + argc);
  for (size_t i = 1; i + 1 < av.size();) {
    const std::string a = av[i];
    if (a == "--m") { krylov = (uint32_t)std::stoul(av[i + 1]); av.erase(av.begin() + i, av.begin() + i + 2); }
    else if (a == "--device") { device = std::stoi(av[i + 1]); av.erase(av.begin() + i, av.begin() + i + 2); }
    else if (a == "--downsample") { ds = (uint32_t)std::stoul(av[i + 1]); av.erase(av.begin() + i, av.begin() + i + 2); }
    else if (a == "--downsample-fft") { ds = (uint32_t)std::stoul(av[i + 1]); ds_fft = true; av.erase(av.begin() + i, av.begin() + i + 2); }
    else if (a == "--diag") { diag_file = av[i + 1]; av.erase(av.begin() + i, av.begin() + i + 2); }
    else if (a == "--devices") { ndev = std::stoi(av[i + 1]); av.erase(av.begin() + i, av.begin() + i + 2); }
    else ++i;
  }
  const int ac = (int)av.size();
  const int nbox = s.dim == 3 ? 6 : 4, nfiles = is_nlse(s) ? 2 : 4;
  const int base = 1 + nbox + s.nscal + nfiles + 3;
  if (s.fields_req ? ac != base + 2 : (ac != base && ac != base + 1)) return usage(s, av[0]);

  int p = 1;
  const uint32_t nx = std::stoul(av[p++]), ny = std::stoul(av[p++]);
  const uint32_t nz = s.dim == 3 ? std::stoul(av[p++]) : 1;
  const double Lx = std::stod(av[p++]), Ly = std::stod(av[p++]);
  if (s.dim == 3) (void)std::stod(av[p++]); // Lz is parsed but dz is computed from Ly (nlse_cubic_driver_3d.cpp:51)
  double sc[2] = {0, 0};
  for (int k = 0; k < s.nscal; ++k) sc[k] = std::stod(av[p++]);
  const std::string in_u = av[p++];
  const std::string in_v = is_nlse(s) ? "" : av[p++];
  const std::string out_u = av[p++];
  const std::string out_v = is_nlse(s) ? "" : av[p++];
  const double T = std::stod(av[p++]);
  const uint32_t nt = std::stoul(av[p++]), ns = std::stoul(av[p++]);
  const std::string m_file = p < ac ? av[p++] : "";
  const std::string c_file = p < ac ? av[p++] : "";
  if (ns == 0 || ns > nt) {
    std::cerr << "Error: need 0 < num_snapshots <= nt (the reference divides by nt/num_snapshots)\n";
    return 1;
  }
  nls::Grid g;
  g.nx = nx; g.ny = ny; g.nz = nz;
  g.dx = 2 * Lx / (nx - 1);
  g.dy = 2 * Ly / (ny - 1);
  g.dz = s.dim == 3 ? 2 * Ly / (nz - 1) : 0.0;
  g.op_sign = s.op_sign;
  const double dt = T / nt;
  const uint32_t freq = nt / ns;
  const size_t n = g.n();
  std::vector<uint64_t> want = s.dim == 3 ? std::vector<uint64_t>{nz, ny, nx} : std::vector<uint64_t>{ny, nx};
  if (ds_fft && s.dim == 3) {
    std::cerr << "Error: --downsample-fft is 2D only (the launcher's downsample_fft is)\n";
    return 1;
  }
  if (ds == 1) {
    std::cerr << "Error: --downsample needs N >= 2\n";
    return 1;
  }
  if (ndev < 1 || device < 0) {
    std::cerr << "Error: need --device >= 0 and --devices >= 1\n";
    return 1;
  }
  std::vector<int> devices;
  for (int k = 0; k < ndev; ++k) devices.push_back(device + k);
  uint64_t B = 1;

  std::vector<double> m(n, s.m_default), c;
  if (!m_file.empty()) {
    std::vector<double> mm;
    bool ok;
    try { ok = load_checked(m_file, want, mm, "m(x)"); }
    catch (const std::exception &e) { std::cerr << "Error loading m: " << e.what() << "\n"; throw std::runtime_error("Aborting"); }
    if (ok) m.swap(mm);
    else if (s.fields_req) throw std::runtime_error("Faulty m");
    else { std::cerr << "Using default m=1.0 everywhere\n"; m.assign(n, 1.0); } // nlse_cubic_driver.cpp:69-74
  }
  if (!c_file.empty()) {
    bool ok;
    try { ok = load_checked(c_file, want, c, "c(x)"); }
    catch (const std::exception &e) { std::cerr << "Error loading c: " << e.what() << "\n"; throw std::runtime_error("Faulty c"); }
    if (!ok) throw std::runtime_error("Faulty c");
    g.c = c.data();
  }

  const auto t0 = std::chrono::steady_clock::now();
  if (is_nlse(s)) {
    if (!run_ensemble<cplx>(s, g, krylov, dt, sc, nt, ns, freq, want, m, in_u, "", out_u, "", devices, B, diag_file, ds, ds_fft)) return 1;
  } else {
    if (!run_ensemble<double>(s, g, krylov, dt, sc, nt, ns, freq, want, m, in_u, in_v, out_u, out_v, devices, B, diag_file, ds, ds_fft)) return 1;
  }
  if (std::getenv("NLS_VERBOSE")) {
    const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    std::cerr << s.name << " [" << s.ref << "]: " << (nt - 1) << " steps of " << n << " sites x " << B
              << " trajectories in " << sec << " s (" << (double)n * B * (nt - 1) / sec
              << " site-steps/s incl. file I/O and staging), krylov m=" << krylov << "\n";
  }
  return 0;
}

} // namespace

int main(int argc, char **argv) {
  std::string self = basename(argv[0]);
  const Spec *spec = nullptr;
  for (const Spec &s : kSpecs)
    if (self == s.name) spec = &s;
  if (!spec && argc >= 2) { // nls_driver <name> args...
    for (const Spec &s : kSpecs)
      if (std::string(argv[1]) == s.name) spec = &s;
    if (spec) { ++argv; --argc; }
  }
  if (!spec) {
    std::cerr << "nls_driver: unknown driver '" << self << "'. Known drivers:\n";
    for (const Spec &s : kSpecs) std::cerr << "  " << s.name << "   (" << s.ref << ")\n";
    return 1;
  }
  return run(*spec, argc, argv); // exceptions (unreadable files, CUDA errors) abort like the reference
}
