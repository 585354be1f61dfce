// Exercises include/block_diag_ilu_b200.hpp the way tests/test_block_diag.cpp of the reference exercises
// block_diag_ilu.hpp: same matrices, same public fields, same expected values (golden vectors of SURVEY 8c).
// Exit code: 0 = all checks passed, 1 = a check failed, 3 = no CUDA device (the compute calls threw, as they must).
#include <cmath>
#include <cstdio>
#include <cstring>
#include <limits>

#include "block_diag_ilu_b200.hpp"

using block_diag_ilu::BlockDiagMatrix;
using block_diag_ilu::ILU;
using block_diag_ilu::ILU_inplace;

static int g_fail = 0;
#define CHECK(cond)                                                            \
    do {                                                                       \
        if (!(cond)) {                                                         \
            std::printf("FAIL %s:%d  %s\n", __FILE__, __LINE__, #cond);        \
            ++g_fail;                                                          \
        }                                                                      \
    } while (0)
static bool close(double a, double b, double tol) { return std::fabs(a - b) <= tol; }

// 2x2 blocks, 3 of them, ndiag = 1 (tests/test_block_diag.cpp:132-163, "m2")
static void fill_m2(BlockDiagMatrix<double> &m)
{
    const double blocks[3][2][2] = {{{5, 3}, {5, 8}}, {{8, 4}, {4, 4}}, {{6, 9}, {2, 7}}};
    for (int b = 0; b < 3; ++b)
        for (int r = 0; r < 2; ++r)
            for (int c = 0; c < 2; ++c) m.block(b, r, c) = blocks[b][r][c];
    const double sub[2][2] = {{1, 2}, {3, 4}}, sup[2][2] = {{2, 3}, {4, 5}};
    for (int b = 0; b < 2; ++b)
        for (int c = 0; c < 2; ++c) {
            m.sub(0, b, c) = sub[b][c];
            m.sup(0, b, c) = sup[b][c];
        }
}

// 6x6 with pivoting blocks and two diagonals (tests/test_block_diag.cpp:165-219, "m4")
static void fill_m4(BlockDiagMatrix<double> &m)
{
    const double A[6][6] = {{-17, 63, .2, 0, .02, 0}, {37, 13, 0, .3, 0, .03}, {.1, 0, 11, 72, -.1, 0},
                            {0, .2, -42, 24, 0, .2},  {.03, 0, -.1, 0, 72, -13}, {0, -.1, 0, .08, 14, -57}};
    for (int r = 0; r < 6; ++r)
        for (int c = 0; c < 6; ++c)
            if (m.valid_index(r, c)) m(r, c) = A[r][c];
}

int main()
{
    // --- container behaviour (host only) ---
    {
        BlockDiagMatrix<double> m(nullptr, 3, 2, 1);
        CHECK(m.m_ld == 8 && m.m_nr == 6 && m.m_nc == 6 && m.m_ndata == (3 * 2 + 2 * 2) * 8);
        CHECK(m.valid_index(2, 0) && !m.valid_index(4, 0) && !m.valid_index(2, 1));
        bool threw = false;
        try { m(4, 0) = 1.0; } catch (const std::runtime_error &e) { threw = std::strcmp(e.what(), "Illegal index") == 0; }
        CHECK(threw);
        threw = false;
        double raw[8];
        try { BlockDiagMatrix<double> v(raw, 1, 2); } catch (const std::runtime_error &) { threw = true; }
        CHECK(threw);
        int piv[5] = {3, 4, 5, 4, 5}, rbc[5], cbr[5];  // tests/test_block_diag.cpp:110-130
        block_diag_ilu::rowpiv2rowbycol(5, piv, rbc);
        block_diag_ilu::rowbycol2colbyrow(5, rbc, cbr);
        const int rbc_ref[5] = {2, 3, 4, 1, 0}, cbr_ref[5] = {4, 3, 0, 1, 2};
        CHECK(!std::memcmp(rbc, rbc_ref, sizeof rbc) && !std::memcmp(cbr, cbr_ref, sizeof cbr));
    }
    try {
        // --- m2: ILU_inplace on a view, peeking at the factors (:132-163) ---
        BlockDiagMatrix<double> m2(nullptr, 3, 2, 1, 0, 2);
        fill_m2(m2);
        {
            double v[6] = {3, 2, 6, 2, 5, 4}, out[6];  // dot_vec, :394-415
            m2.dot_vec(v, out);
            const double ref[6] = {5 * 3 + 3 * 2 + 2 * 6, 5 * 3 + 8 * 2 + 3 * 2, 1 * 3 + 8 * 6 + 4 * 2 + 4 * 5,
                                   2 * 2 + 4 * 6 + 4 * 2 + 5 * 4, 3 * 6 + 6 * 5 + 9 * 4, 4 * 2 + 2 * 5 + 7 * 4};
            for (int i = 0; i < 6; ++i) CHECK(close(out[i], ref[i], 1e-13));
        }
        ILU_inplace<double> ilu(&m2);
        CHECK(ilu.m_view == &m2);
        CHECK(m2.sub(0, 0, 0) == 1 / 5. && m2.sub(0, 0, 1) == 2 / 5. && m2.sub(0, 1, 0) == 3 / 8. && m2.sub(0, 1, 1) == 4 / 2.);
        CHECK(m2.sup(0, 0, 0) == 2 && m2.sup(0, 0, 1) == 3 && m2.sup(0, 1, 0) == 4 && m2.sup(0, 1, 1) == 5);
        CHECK(m2.block(0, 0, 0) == 5 && m2.block(0, 1, 0) == 1 && m2.block(0, 1, 1) == 5);
        for (int i = 0; i < 6; ++i) CHECK(ilu.m_ipiv[i] == (i % 2) + 1 && ilu.m_rowbycol[i] == i % 2 && ilu.m_colbyrow[i] == i % 2);
        const double b[6] = {65, 202, 11, 65, 60, 121};
        const double xref[6] = {-31.47775, 53.42125, 31.0625, -43.36875, -19.25625, 19.5875};
        double x[6];
        CHECK(ilu.solve(b, x) == 0);
        for (int i = 0; i < 6; ++i) CHECK(close(x[i], xref[i], 1e-13));
        double bn[6] = {65, 202, std::numeric_limits<double>::quiet_NaN(), 65, 60, 121};
        CHECK(ilu.solve(bn, x) == 3);  // 1-based index of the first NaN (:500-501)

        // --- m4: owning ILU, pivoting, two diagonals (:165-219) ---
        BlockDiagMatrix<double> m4(nullptr, 3, 2, 2, 0, 2);
        fill_m4(m4);
        ILU<double> ilu4(m4);
        CHECK(m4.block(0, 0, 0) == -17);  // the source is untouched: ILU factors its own copy (:546-547)
        const int rbc4[6] = {1, 0, 1, 0, 0, 1};
        for (int i = 0; i < 6; ++i) CHECK(ilu4.m_ilu_inplace.m_rowbycol[i] == rbc4[i] && ilu4.m_ilu_inplace.m_colbyrow[i] == rbc4[i]);
        const double a_ = 63 + 17 / 37. * 13, b_ = 72 + 11 / 42. * 24;
        const BlockDiagMatrix<double> &f = *ilu4.m_ilu_inplace.m_view;
        CHECK(close(f.sub(0, 0, 0), .1 / 37, 1e-15) && close(f.sub(0, 0, 1), .2 / a_, 1e-15) && close(f.sub(0, 1, 0), .1 / 42, 1e-15) &&
              close(f.sub(0, 1, 1), .08 / b_, 1e-15) && close(f.sub(1, 0, 0), .03 / 37, 1e-15) && close(f.sub(1, 0, 1), -.1 / a_, 1e-15));
        const double b4[6] = {-62, 207, 11, -14, 25, -167};
        const double x4[6] = {5.42317616680147374, 0.478588898186963929, 0.400565700557765081,
                              0.0873749367816923223, 0.914791409109598774, 3.15378902934640371};
        CHECK(ilu4.solve(b4, x) == 0);
        for (int i = 0; i < 6; ++i) CHECK(close(x[i], x4[i], 1e-14));

        // --- set_to_eye_plus_scaled_mtx (:288-290) and a singular block (:483-484) ---
        BlockDiagMatrix<double> m5(nullptr, 3, 2, 2, 0, 2);
        m5.set_to_eye_plus_scaled_mtx(-0.5, m4);
        CHECK(m5.block(0, 0, 0) == 1 + 17 * 0.5 && m5.block(0, 1, 0) == -18.5 && m5.sub(1, 0, 1) == 0.05);
        BlockDiagMatrix<double> sing(nullptr, 2, 2, 1, 0, 2);
        sing.block(0, 0, 0) = 1; sing.block(0, 1, 1) = 1;
        sing.block(1, 0, 0) = 1; sing.block(1, 0, 1) = 2; sing.block(1, 1, 0) = 2; sing.block(1, 1, 1) = 4;
        bool threw = false;
        try { ILU_inplace<double> bad(&sing); } catch (const std::runtime_error &e) { threw = std::strcmp(e.what(), "ILU failed!") == 0; }
        CHECK(threw);

        // --- MatrixBase source constructor (block_diag_ilu.hpp:124-159): entries inside the pattern are copied ---
        {
            struct Dense6 {
                double a[6][6];
                double operator()(int r, int c) const { return a[r][c]; }
            } src;
            for (int r = 0; r < 6; ++r)
                for (int c = 0; c < 6; ++c) src.a[r][c] = 10 * (r + 1) + (c + 1);
            BlockDiagMatrix<double> fromsrc(src, 3, 2, 1, 1);
            CHECK(fromsrc.m_ld == 8 && fromsrc.m_nsat == 1);
            for (int r = 0; r < 6; ++r)
                for (int c = 0; c < 6; ++c)
                    if (fromsrc.valid_index(r, c)) CHECK(fromsrc(r, c) == src.a[r][c]);
            CHECK(fromsrc.bot(0, 0, 1) == src.a[5][1] && fromsrc.top(0, 0, 0) == src.a[0][4]);
            BlockDiagMatrix<double> again(m4, 3, 2, 2, 0, 2);  // another BlockDiagMatrix as the source (:291-293)
            for (int i = 0; i < m4.m_ndata; ++i) CHECK(again.m_data[i] == m4.m_data[i]);
        }

        // --- Real_t = float: the m2 case in single precision (the template parameter of hpp:104, 426) ---
        {
            BlockDiagMatrix<float> f2(nullptr, 3, 2, 1);
            CHECK(f2.m_ld == 16);  // 64-byte rows: 16 floats (hpp:98)
            const float blocks[3][2][2] = {{{5, 3}, {5, 8}}, {{8, 4}, {4, 4}}, {{6, 9}, {2, 7}}};
            const float subv[2][2] = {{1, 2}, {3, 4}}, supv[2][2] = {{2, 3}, {4, 5}};
            for (int bq = 0; bq < 3; ++bq)
                for (int r = 0; r < 2; ++r)
                    for (int c = 0; c < 2; ++c) f2.block(bq, r, c) = blocks[bq][r][c];
            for (int bq = 0; bq < 2; ++bq)
                for (int c = 0; c < 2; ++c) { f2.sub(0, bq, c) = subv[bq][c]; f2.sup(0, bq, c) = supv[bq][c]; }
            float vf[6] = {3, 2, 6, 2, 5, 4}, of[6];
            f2.dot_vec(vf, of);
            CHECK(of[0] == 33.f && of[1] == 37.f && of[5] == 46.f);
            ILU<float> iluf(f2);
            const float bf[6] = {65, 202, 11, 65, 60, 121};
            float xf[6];
            CHECK(iluf.solve(bf, xf) == 0);
            for (int i = 0; i < 6; ++i) CHECK(close(xf[i], xref[i], 2e-4));
            CHECK(std::fabs(iluf.m_ilu_inplace.m_view->sub(0, 1, 0) - 3 / 8.f) < 1e-7f);
            float bnf[6] = {65, 202, 11, std::numeric_limits<float>::quiet_NaN(), 60, 121};
            CHECK(iluf.solve(bnf, xf) == 4);
        }
    } catch (const std::runtime_error &e) {
        if (std::strstr(e.what(), "no CUDA device")) {
            std::printf("no CUDA device: %s\n", e.what());
            return g_fail ? 1 : 3;
        }
        std::printf("unexpected exception: %s\n", e.what());
        return 1;
    }
    std::printf(g_fail ? "%d checks failed\n" : "shim ok\n", g_fail);
    return g_fail ? 1 : 0;
}
