// Host-side test of include/g1mpc_mpc.hpp, the header-only g1_mpc::MPC adapter over the C-ABI.
//   test_mpc_adapter in.bin out.bin    (GPU box)  -- reads one problem (x0[13] | x_ref[N,13] | contact[N,4] | c_horizon[N,12] |
//       p_com[N,3], N = 10, raw doubles), runs the call sequence of the reference node (mpc_to_wbid_node.cpp:392-394,455-494:
//       setX0, setXRefHor, updateMPC(contact, c_horizon, p_com_horizon, &x_current, false)), then the getters the node reads
//       (:329-330,360,478), then updateMPC(..., one_rollout = true), solveQP() and computeRollout(); writes every result to out.bin.
//   test_mpc_adapter --no-gpu          (build container) -- the constructor must throw "no CUDA device ... no CPU fallback".
// tests/test_cpp_adapter.py drives both and compares out.bin with the Python path.
#include <cstdio>
#include <cstring>
#include <vector>

#include "g1mpc_mpc.hpp"

static int fail(const char *m) { std::fprintf(stderr, "FAIL: %s\n", m); return 1; }

int main(int argc, char **argv)
{
    if (argc == 2 && std::strcmp(argv[1], "--no-gpu") == 0) {
        try {
            g1_mpc::MPC m;
            return fail("constructor succeeded without a CUDA device");
        } catch (const std::runtime_error &e) {
            return std::strstr(e.what(), "no CPU fallback") ? 0 : fail(e.what());
        }
    }
    if (argc != 3) return fail("usage: test_mpc_adapter in.bin out.bin | --no-gpu");
    const int N = 10, NC = 4;
    std::vector<double> in(13 + N * 13 + N * NC + N * 12 + N * 3);
    FILE *f = std::fopen(argv[1], "rb");
    if (!f || std::fread(in.data(), sizeof(double), in.size(), f) != in.size()) return fail("cannot read the input file");
    std::fclose(f);
    g1_mpc::MPC mpc;                                         // defaults of mpc.hpp:22: mu 0.8, fz 10..1000, dt 0.05, N 10
    if (g1_mpc::MPC::horizon_length_ != N || mpc.getNumContacts() != NC || mpc.getDt() != 0.05) return fail("defaults");
    g1_mpc::Vector x0(13);
    g1_mpc::Matrix xr(N, 13), ct(N, NC), ch(N, 12), pc(N, 3);
    const double *p = in.data();
    for (int i = 0; i < 13; i++) x0(i) = *p++;
    for (int i = 0; i < N * 13; i++) xr.data()[i] = *p++;
    for (int i = 0; i < N * NC; i++) ct.data()[i] = *p++;
    for (int i = 0; i < N * 12; i++) ch.data()[i] = *p++;
    for (int i = 0; i < N * 3; i++) pc.data()[i] = *p++;
    mpc.setX0(x0);
    mpc.setXRefHor(xr);
    std::vector<double> out;
    auto put = [&out](const double *d, int n) { out.insert(out.end(), d, d + n); };
    // (1) the node's call
    auto res = mpc.updateMPC(ct, ch, pc, &x0, false);
    if (mpc.status() != G1MPC_SOLVED) return fail("status");
    put(res.first.data(), N * 12); put(res.second.data(), N * 13);
    put(mpc.getUOpt0().data(), 12); put(mpc.getr().data(), 3);
    if (mpc.getXOpt().rows() != N || mpc.getUOpt().cols() != 12 || mpc.getContactHorizon()(0, 0) != ct(0, 0) ||
        mpc.getCHorizon()(N - 1, 11) != ch(N - 1, 11) || mpc.getPComHorizon()(3, 1) != pc(3, 1) || mpc.getX0()(5) != x0(5) ||
        mpc.getXRefHor()(2, 3) != xr(2, 3)) return fail("getters");
    // (2) the WBID-side call pattern: one rollout step only (mpc_to_wbid.py:158)
    g1_mpc::Matrix zero(N, 13);
    mpc.setXOpt(zero);
    auto res1 = mpc.updateMPC(ct, ch, pc, &x0, true);
    put(res1.second.data(), N * 13);
    // (3) main.py's sequence: step methods, solveQP, then computeRollout of the stored u_opt (main.py:290-310)
    mpc.extractPsi(); mpc.calculateRotationMatrixT(); mpc.setQ(); mpc.setR(); mpc.calculateAContinuous(); mpc.calculateADiscrete();
    mpc.calculateBContinuous(ch, pc); mpc.calculateBDiscrete(); mpc.calculateAqp(); mpc.calculateBqp(); mpc.calculateAc();
    mpc.calculateBounds(ct); mpc.calculateHessian(); mpc.calculateGradient();
    g1_mpc::Vector sol = mpc.solveQP();
    put(sol.data(), N * 12);
    mpc.setXOpt(zero);
    mpc.computeRollout(&x0, false);
    put(mpc.getXOpt().data(), N * 13);
    // (4) a caller-modified u_opt is what computeRollout rolls out (mpc.cpp:427-446 reads the member)
    g1_mpc::Matrix u2 = mpc.getUOpt();
    for (int i = 0; i < N * 12; i++) u2.data()[i] *= 0.5;
    mpc.setUOpt(u2);
    mpc.computeRollout(&x0, false);
    put(mpc.getXOpt().data(), N * 13);
    f = std::fopen(argv[2], "wb");
    if (!f || std::fwrite(out.data(), sizeof(double), out.size(), f) != out.size()) return fail("cannot write the output file");
    std::fclose(f);
    std::printf("ok: %zu doubles, %d iterations\n", out.size(), mpc.iterations());
    return 0;
}
