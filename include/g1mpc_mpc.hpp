// g1mpc_mpc.hpp -- header-only C++ adapter: the public surface of the reference class g1_mpc::MPC
// (g1_mpc/include/mpc.hpp:15-133) over the C-ABI of g1mpc_b200.h.  No Eigen, no CUDA headers: link with
// -lg1mpc_b200 only.  A maintainer of the reference node swaps `#include "mpc.hpp"` for this file (or keeps Eigen at
// the call site and converts through Matrix::data()).
//
//   * g1_mpc::Matrix / g1_mpc::Vector: minimal dense containers with Eigen's accessors used by the node
//     (rows(), cols(), operator()(i,j), operator()(i), size(), data(), setZero(), row-major storage because the C-ABI is
//     row-major; Eigen's default is column-major -- convert with Eigen::Map<RowMajor> at the boundary).
//   * Constructor, initMatrices, updateMPC, solveQP, computeRollout, extractPsi, every getter/setter of
//     mpc.hpp:90-130, the statics g_ and horizon_length_.
//   * The fifteen calculate*/set* step methods exist and keep their signatures; the matrices they fill in the reference
//     are private members without getters (mpc.hpp:133-199), so here they only record their arguments: the fused kernel
//     rebuilds everything from the inputs on the next solveQP()/updateMPC() (g1mpc_emit_qp_batch materialises the
//     matrices for callers that want to look at them).
//   * Error behaviour: like the reference, numerical trouble never throws (status()/iterations() report it; the
//     reference ignores the solver status, mpc.cpp:405); misuse and CUDA failures throw std::runtime_error with
//     g1mpc_last_error().
#pragma once

#include <cmath>
#include <cstdint>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "g1mpc_b200.h"

namespace g1_mpc
{

class Vector
{
public:
    Vector() {}
    explicit Vector(int n) : v_(static_cast<size_t>(n), 0.0) {}
    Vector(std::initializer_list<double> l) : v_(l) {}
    int size() const { return static_cast<int>(v_.size()); }
    int rows() const { return size(); }
    double &operator()(int i) { return v_[static_cast<size_t>(i)]; }
    double operator()(int i) const { return v_[static_cast<size_t>(i)]; }
    double &operator[](int i) { return v_[static_cast<size_t>(i)]; }
    double operator[](int i) const { return v_[static_cast<size_t>(i)]; }
    double *data() { return v_.data(); }
    const double *data() const { return v_.data(); }
    void setZero() { v_.assign(v_.size(), 0.0); }
    void resize(int n) { v_.assign(static_cast<size_t>(n), 0.0); }
private:
    std::vector<double> v_;
};

class Matrix
{
public:
    Matrix() : r_(0), c_(0) {}
    Matrix(int r, int c) : r_(r), c_(c), v_(static_cast<size_t>(r) * c, 0.0) {}
    static Matrix Zero(int r, int c) { return Matrix(r, c); }
    int rows() const { return r_; }
    int cols() const { return c_; }
    double &operator()(int i, int j) { return v_[static_cast<size_t>(i) * c_ + j]; }
    double operator()(int i, int j) const { return v_[static_cast<size_t>(i) * c_ + j]; }
    double *data() { return v_.data(); }               // row-major
    const double *data() const { return v_.data(); }
    void setZero() { v_.assign(v_.size(), 0.0); }
    void resize(int r, int c) { r_ = r; c_ = c; v_.assign(static_cast<size_t>(r) * c, 0.0); }
    Vector row(int i) const { Vector o(c_); for (int j = 0; j < c_; j++) o(j) = (*this)(i, j); return o; }
private:
    int r_, c_;
    std::vector<double> v_;
};

class MPC
{
public:
    static constexpr double g_ = -9.80665;              // mpc.cpp:7
    static inline int horizon_length_ = 10;             // process-wide static in the reference (mpc.hpp:19, mpc.cpp:8,19)

    // mpc.hpp:22 / mpc.cpp:10-103
    MPC(double mu = 0.8, double fz_min = 10.0, double fz_max = 1000.0, double dt = 0.05, int horizon_length = 10, int device = 0)
        : dt_(dt), num_contacts_(4), num_states_(13), num_controls_(12), mu_(mu), fz_min_(fz_min), fz_max_(fz_max), psi_(0.0),
          device_(device)
    {
        horizon_length_ = horizon_length;
        n_ = horizon_length;
        initMatrices();
        g1mpc_params p{};
        p.horizon = n_; p.num_contacts = num_contacts_;
        p.mu = mu_; p.fz_min = fz_min_; p.fz_max = fz_max_; p.dt = dt_;
        p.mass = robot_mass_; p.gravity = g_;
        const double I[9] = {8.20564e-2, 0, 0, 0, 8.05015e-2, 0, 0, 0, 0.32353e-2};                     // mpc.cpp:25-27
        const double q[13] = {8e2, 75, 125e1, 3e4, 8e5, 8e5, 2e3, 4e3, 3e2, 5e2, 8e3, 5e2, 0};          // mpc.cpp:90-96
        for (int i = 0; i < 9; i++) p.inertia_body[i] = I[i];
        for (int i = 0; i < 13; i++) p.q_weights[i] = q[i];
        p.r_weight = 1e-3;                                                                              // mpc.cpp:98
        p.cost_scale = 2.0;                                                                             // mpc.cpp:361,375
        if (g1mpc_create(&p, device_, &h_) != G1MPC_OK) throw std::runtime_error(std::string("g1mpc_create: ") + g1mpc_last_error());
    }
    ~MPC() { g1mpc_destroy(h_); }
    MPC(const MPC &) = delete;
    MPC &operator=(const MPC &) = delete;

    // mpc.cpp:110-151
    void initMatrices()
    {
        x0_ = Vector(num_states_);
        x_ref_hor_ = Matrix(n_, num_states_);
        u_opt_ = Matrix(n_, num_controls_);
        x_opt_ = Matrix(n_, num_states_);
        u_opt0_ = Vector(num_controls_);
        r_ = Vector(3);
        contact_horizon_ = Matrix(n_, num_contacts_);
        c_horizon_ = Matrix(n_, 3 * num_contacts_);
        p_com_horizon_ = Matrix(n_, 3);
    }

    // mpc.cpp:153-162
    double extractPsi()
    {
        double s = 0.0;
        for (int i = 0; i < n_; i++) s += x_ref_hor_(i, 2);
        psi_ = s / n_;
        return psi_;
    }
    // mpc.cpp:164-171
    Matrix vectorToSkewSymmetricMatrix(const Vector &v) const
    {
        Matrix m(3, 3);
        m(0, 1) = -v(2); m(0, 2) = v(1); m(1, 0) = v(2); m(1, 2) = -v(0); m(2, 0) = -v(1); m(2, 1) = v(0);
        return m;
    }
    // The step methods of mpc.hpp:37-74: their products are private in the reference; the kernel rebuilds them (see header comment).
    void calculateRotationMatrixT() {}
    void setQ() {}
    void setR() {}
    void calculateAContinuous() {}
    void calculateADiscrete() {}
    void calculateBContinuous(const Matrix &c_horizon, const Matrix &p_com_horizon)
    {
        check(c_horizon, n_, 3 * num_contacts_, "c_horizon"); check(p_com_horizon, n_, 3, "p_com_horizon");
        c_horizon_ = c_horizon; p_com_horizon_ = p_com_horizon;
        for (int a = 0; a < 3; a++) r_(a) = c_horizon_(n_ - 1, 3 * (num_contacts_ - 1) + a) - p_com_horizon_(n_ - 1, a);   // mpc.cpp:229-233, last pass
    }
    void calculateBDiscrete() {}
    void calculateAqp() {}
    void calculateBqp() {}
    void calculateAc() {}
    void calculateBounds(const Matrix &contact) { check(contact, n_, num_contacts_, "contact"); contact_horizon_ = contact; }
    void calculateHessian() {}
    void calculateGradient() {}

    // mpc.cpp:378-425: solves the QP of the current inputs; sets u_opt_, u_opt0_ and returns the decision vector
    Vector solveQP()
    {
        Matrix scratch = x_opt_;
        run(nullptr, true, scratch);
        Vector x(n_ * num_controls_);
        for (int i = 0; i < n_ * num_controls_; i++) x(i) = u_opt_.data()[i];
        return x;
    }

    // mpc.cpp:448-473
    std::pair<Matrix, Matrix> updateMPC(const Matrix &contact, const Matrix &c_horizon, const Matrix &p_com_horizon,
                                        const Vector *x_current = nullptr, bool one_rollout = false)
    {
        extractPsi();
        calculateBContinuous(c_horizon, p_com_horizon);
        calculateBounds(contact);
        if (x_current && x_current->size() != num_states_) throw std::runtime_error("updateMPC: x_current must have 13 entries");
        run(x_current ? x_current->data() : nullptr, one_rollout, x_opt_);
        return {u_opt_, x_opt_};                        // by value, as mpc.cpp:472
    }

    // mpc.cpp:427-446: rollout of the CURRENT u_opt_
    void computeRollout(const Vector *x_current = nullptr, bool only_first_step = false)
    {
        if (g1mpc_rollout_host(h_, 1, x_ref_hor_.data(), c_horizon_.data(), p_com_horizon_.data(), u_opt_.data(),
                               x_current ? x_current->data() : nullptr, only_first_step ? 1 : 0, x_opt_.data()) != G1MPC_OK)
            throw std::runtime_error(std::string("g1mpc_rollout_host: ") + g1mpc_last_error());
    }

    // mpc.hpp:90-130
    const Vector &getX0() const { return x0_; }
    void setX0(const Vector &x0) { x0_ = x0; }
    const Matrix &getXRefHor() const { return x_ref_hor_; }
    void setXRefHor(const Matrix &x_ref_hor) { x_ref_hor_ = x_ref_hor; }
    const Matrix &getUOpt() const { return u_opt_; }
    void setUOpt(const Matrix &u_opt) { u_opt_ = u_opt; }
    const Matrix &getXOpt() const { return x_opt_; }
    void setXOpt(const Matrix &x_opt) { x_opt_ = x_opt; }
    const Vector &getUOpt0() const { return u_opt0_; }
    void setUOpt0(const Vector &u_opt0) { u_opt0_ = u_opt0; }
    const Vector &getr() const { return r_; }
    void setr(const Vector &r) { r_ = r; }
    const Matrix &getContactHorizon() const { return contact_horizon_; }
    void setContactHorizon(const Matrix &contact_horizon) { contact_horizon_ = contact_horizon; }
    const Matrix &getPComHorizon() const { return p_com_horizon_; }
    void setPComHorizon(const Matrix &p_com_horizon) { p_com_horizon_ = p_com_horizon; }
    const Matrix &getCHorizon() const { return c_horizon_; }
    void setCHorizon(const Matrix &c_horizon) { c_horizon_ = c_horizon; }
    int getNumContacts() const { return num_contacts_; }
    double getDt() const { return dt_; }
    void debugPrintMatrixDimensions() const {}

    // additions of this implementation (the reference discards the solver status, mpc.cpp:405)
    int status() const { return status_; }
    int iterations() const { return iters_; }
    void setWarmStart(bool on) { g1mpc_set_host_warm_start(h_, on ? 1 : 0, 1); }   // what mpc.cpp:386 asks OSQP for
    g1mpc_handle *handle() const { return h_; }

private:
    static void check(const Matrix &m, int r, int c, const char *name)
    {
        if (m.rows() != r || m.cols() != c)
            throw std::runtime_error(std::string(name) + ": expected " + std::to_string(r) + " x " + std::to_string(c));
    }
    void run(const double *x_current, bool one_rollout, Matrix &x_opt)
    {
        if (x0_.size() != num_states_) throw std::runtime_error("x0 must have 13 entries");
        check(x_ref_hor_, n_, num_states_, "x_ref_hor");
        int32_t st = 0, it = 0;
        const int rc = g1mpc_update_host(h_, 1, x0_.data(), x_ref_hor_.data(), contact_horizon_.data(), c_horizon_.data(),
                                         p_com_horizon_.data(), x_current, one_rollout ? 1 : 0, u_opt_.data(), x_opt.data(), &st, &it,
                                         nullptr);
        if (rc != G1MPC_OK) throw std::runtime_error(std::string("g1mpc_update_host: ") + g1mpc_last_error());
        status_ = st; iters_ = it;
        for (int j = 0; j < num_controls_; j++) u_opt0_(j) = u_opt_(0, j);       // mpc.cpp:421-422
    }

    const double robot_mass_ = 34.13385728;             // mpc.hpp:136
    double dt_;
    int num_contacts_, num_states_, num_controls_;
    const int num_bounds_ = 5;
    double mu_, fz_min_, fz_max_, psi_;
    int device_, n_ = 10, status_ = 0, iters_ = 0;
    g1mpc_handle *h_ = nullptr;
    Vector x0_, u_opt0_, r_;
    Matrix x_ref_hor_, u_opt_, x_opt_, contact_horizon_, c_horizon_, p_com_horizon_;
};

} // namespace g1_mpc
