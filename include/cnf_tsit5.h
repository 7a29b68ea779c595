/* cnf_tsit5.h — every numerical constant of the adaptive Tsit5 solve, in ONE place.
 *
 * The reference calls `SciMLBase.solve(prob; icnf.sol_kwargs...)` with no algorithm
 * (reference src/base_icnf.jl:141), i.e. OrdinaryDiffEqDefault, which starts on Tsit5.
 * That package is a third-party, un-vendored dependency (Project.toml:41-70), so the
 * constants below restate the published Tsitouras 5(4) tableau (Tsitouras 2011) and the
 * OrdinaryDiffEq defaults documented in SURVEY.md Appendix A.  PARITY UNPINNED: no golden
 * vector of the reference exists for them.
 *
 * Shared by the CUDA library (continuousnf.jl_b200/csrc) and by the CPU oracle (oracle/),
 * which is test infrastructure only.
 */
#ifndef CNF_TSIT5_H
#define CNF_TSIT5_H

/* Butcher tableau, SURVEY.md A.2 */
#define CNF_A21 0.161
#define CNF_A31 (-0.008480655492356989)
#define CNF_A32 0.335480655492357
#define CNF_A41 2.8971530571054935
#define CNF_A42 (-6.359448489975075)
#define CNF_A43 4.3622954328695815
#define CNF_A51 5.325864828439257
#define CNF_A52 (-11.748883564062828)
#define CNF_A53 7.4955393428898365
#define CNF_A54 (-0.09249506636175525)
#define CNF_A61 5.86145544294642
#define CNF_A62 (-12.92096931784711)
#define CNF_A63 8.159367898576159
#define CNF_A64 (-0.071584973281401)
#define CNF_A65 (-0.028269050394068383)
#define CNF_A71 0.09646076681806523
#define CNF_A72 0.01
#define CNF_A73 0.4798896504144996
#define CNF_A74 1.379008574103742
#define CNF_A75 (-3.290069515436081)
#define CNF_A76 2.324710524099774
/* embedded error weights b~ (sum to 0) */
#define CNF_BT1 (-0.00178001105222577714)
#define CNF_BT2 (-0.0008164344596567469)
#define CNF_BT3 0.007880878010261995
#define CNF_BT4 (-0.1447110071732629)
#define CNF_BT5 0.5823571654525552
#define CNF_BT6 (-0.45808210592918697)
#define CNF_BT7 0.015151515151515152

/* PI controller, SURVEY.md A.5 */
#define CNF_BETA1 (7.0 / 50.0)
#define CNF_BETA2 (2.0 / 25.0)
#define CNF_GAMMA 0.9
#define CNF_QMIN 0.2
#define CNF_QMAX 10.0
#define CNF_QOLDINIT 1e-4

/* auto-switch of the reference's default algorithm, reported only: Tsit5's stability size and the
 * non-stiff tolerance of OrdinaryDiffEq's AutoSwitch; 10 consecutive stiff steps would switch to a stiff method */
#define CNF_TSIT5_STABILITY 3.5068
#define CNF_STIFFTOL 0.9f
#define CNF_MAXSTIFFSTEP 10

/* solver defaults when sol_kwargs is empty, SURVEY.md A.4 */
#define CNF_DEFAULT_RELTOL 1e-3f
#define CNF_DEFAULT_ABSTOL 1e-6f
#define CNF_DEFAULT_MAXITERS 100000LL

/* NNlib.tanh_fast Float32 rational: x*(n(x^2)/d(x^2)), |x|^2<66 else sign(x) */
#define CNF_TANH_N1 0.1346604f
#define CNF_TANH_N2 0.0035974074f
#define CNF_TANH_N3 2.2332108e-5f
#define CNF_TANH_N4 1.587199e-8f
#define CNF_TANH_D1 0.4679937f
#define CNF_TANH_D2 0.026262015f
#define CNF_TANH_D3 0.0003453992f
#define CNF_TANH_D4 8.7767893e-7f
#define CNF_TANH_CLAMP 66.0f

#endif /* CNF_TSIT5_H */
