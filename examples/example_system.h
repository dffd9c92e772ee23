// The 6x6 block-tridiagonal system used by the upstream GBD-PCG examples (nx = 2, N = 3; pendulum LQR Schur
// complement, negative definite) plus its fp64 solution, derived in tests/golden/make_c1_kat.py.
#pragma once
namespace example_system {
constexpr unsigned kStateSize = 2, kKnotPoints = 3;
// block-row k = [left | diag | right], each block column-major; first and last block are padding
template <typename T>
inline void fill(T (&S)[36], T (&gamma)[6], T (&lambda)[6])
{
    const double s[36] = {0, 0, 0, 0,                   -.999, 0, 0, -.999,            .999, .0999, -.98, .999,
                          .999, -.98, .0999, .999,      -2.008, .8801, .8801, -3.0584, .999, .0999, -.98, .999,
                          .999, -.98, .0999, .999,      -1.019, .8801, .8801, -2.0694, 0, 0, 0, 0};
    const double g[6] = {3.1385, 0, 0, 3.0788, .0031, 3.0788};
    for (int i = 0; i < 36; i++) S[i] = static_cast<T>(s[i]);
    for (int i = 0; i < 6; i++) { gamma[i] = static_cast<T>(g[i]); lambda[i] = T(0); }
}
constexpr double kLambdaStar[6] = {-303.7029860857, -46.4159396812, -315.1763026319, -14.8983094180, -298.7908619196, 13.5037826883};
}  // namespace example_system
