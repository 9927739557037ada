"""The .ChF -> Python translator itself (oracle/chf2py.py), on small Chombo-Fortran texts written for this test: the
semantics that the pinning of the oracle rests on -- macro expansion per CH_SPACEDIM, Fortran's DO trip count and exit
value, INTEGER division and mod, block / one-line IF, continuation lines, F90 array sections, cpp conditionals and
macros, by-reference scalar outputs."""
import os
import tempfile

import numpy as np
import pytest

from oracle import chf2py
from oracle.chf_runtime import Box, Fab, FInt, Vec, f_doexit, frange

SRC = '''
#define SCALEIT(x) (two * (x))
C     a comment line
      subroutine TESTKERNEL (
     &     CHF_FRA[phi],
     &     CHF_CONST_FRA1[w],
     &     CHF_BOX[region],
     &     CHF_CONST_REALVECT[dx],
     &     CHF_CONST_INT[dir],
     &     CHF_REAL[total],
     &     CHF_INT[count])

      integer CHF_AUTODECL[i]
      integer CHF_AUTODECL[ii]
      integer n, r, half_n
      REAL_T s

      CHF_AUTOID[ii; dir]
      half_n = (CHF_UBOUND[region; 0] - CHF_LBOUND[region; 0] + 1) / 2   ! integer division
      count = half_n + mod(-7, 3)
      do n = 0, CHF_NCOMP[phi] - 1
        CHF_AUTOMULTIDO[region; i]
          s = w(CHF_OFFSETIX[i; +ii]) - w(CHF_OFFSETIX[i; -ii])
          if (s .lt. zero) s = -s
          if ((s .gt. one) .and. (n .eq. 0)) then
            phi(CHF_AUTOIX[i], n) = SCALEIT(s)
     &                            / dx(dir)
          else if (s .gt. half) then
            phi(CHF_AUTOIX[i], n) = s
          else
            phi(CHF_AUTOIX[i], n) = zero
          endif
          total = total + phi(CHF_AUTOIX[i], n)
        CHF_ENDDO
      enddo
#if CH_SPACEDIM == 3
      count = count + 100
#else
      count = count + 200
#endif
      do r = 1, 3
        count = count + 1
      enddo
      count = count + 1000 * r
      return
      end

      subroutine SECTIONS (
     &     CHF_FRA1[a],
     &     CHF_CONST_FRA1[b],
     &     CHF_BOX[box])
      a(iboxlo0:iboxhi0, iboxlo1:iboxhi1) = b(iboxlo0+1:iboxhi0+1, iboxlo1:iboxhi1) * three
     &                                    - b(iboxlo0:iboxhi0, iboxlo1:iboxhi1)
      return
      end
'''


def _module(D, only=None):
    with tempfile.TemporaryDirectory() as td:
        p = os.path.join(td, "TestF.ChF")
        open(p, "w").write(SRC)
        good, bad = chf2py.translate_file(p, D, only=only)
    assert not (set(bad) - {"sections"} if D == 3 else bad), bad
    ns = {}
    exec("import math\nfrom oracle.chf_runtime import *\n" + "\n".join(good.values()), ns)
    return ns


@pytest.mark.parametrize("D", [2, 3])
def test_translated_kernel_matches_a_hand_evaluation(D):
    ns = _module(D, only=["testkernel"])
    rng = np.random.default_rng(1)
    lo = [3, -2, 1][:D]
    n = [5, 4, 3][:D]
    hi = [lo[d] + n[d] - 1 for d in range(D)]
    w = rng.uniform(-2, 2, size=(1,) + tuple(m + 2 for m in reversed(n)))
    phi = np.zeros((2,) + tuple(reversed(n)))
    dx = [0.5, 0.25, 0.125][:D]
    for direction in range(D):
        phi[...] = 0.0
        total, count = ns["testkernel"](Fab(phi, lo), Fab(w, [x - 1 for x in lo]), Box(lo, hi), Vec(dx), direction, 0.0, 0)
        # by hand with numpy
        ax = D - direction
        core = tuple(slice(1, -1) for _ in range(D))
        up = list(core)
        dn = list(core)
        up[ax - 1] = slice(2, None)
        dn[ax - 1] = slice(0, -2)
        s = np.abs(w[0][tuple(up)] - w[0][tuple(dn)])
        want0 = np.where(s > 1.0, (2.0 * s) / dx[direction], np.where(s > 0.5, s, 0.0))
        want1 = np.where(s > 0.5, s, 0.0)
        assert np.array_equal(phi[0], want0) and np.array_equal(phi[1], want1)
        # accumulation order: component outermost, then the last index, ..., first index innermost
        acc = 0.0
        for arr in (want0, want1):
            for v in arr.reshape(-1):
                acc = acc + v
        assert total == acc
        # (5 / 2 = 2) + mod(-7, 3) (= -1: sign of the dividend) + 100 or 200 + three trips + 1000 * (DO variable after the loop = 4)
        assert count == 2 - 1 + (100 if D == 3 else 200) + 3 + 4000
        assert isinstance(count, FInt)


def test_array_sections_become_loops():
    ns = _module(2, only=["sections"])
    rng = np.random.default_rng(2)
    b = rng.uniform(-1, 1, size=(1, 4, 7))
    a = np.zeros((1, 4, 6))
    ns["sections"](Fab(a, [0, 0]), Fab(b, [0, 0]), Box([0, 0], [5, 3]))
    assert np.array_equal(a[0], b[0][:, 1:7] * 3.0 - b[0][:, 0:6])


def test_fortran_integer_and_do_semantics():
    assert FInt(7) / 2 == 3 and FInt(-7) / 2 == -3 and 7 / FInt(-2) == -3          # truncation toward zero
    assert isinstance(FInt(7) / 2, FInt) and (FInt(7) / 2.0) == 3.5
    assert list(frange(1, 3)) == [1, 2, 3] and list(frange(5, 1, -2)) == [5, 3, 1] and list(frange(2, 1)) == []
    assert f_doexit(1, 3) == 4 and f_doexit(5, 1, -2) == -1 and f_doexit(2, 1) == 2


def test_unknown_constructs_are_refused_not_guessed():
    bad_src = SRC.replace("      return\n      end\n\n      subroutine SECTIONS", "      call SOMETHINGELSE(count)\n      return\n      end\n\n      subroutine SECTIONS", 1)
    with tempfile.TemporaryDirectory() as td:
        p = os.path.join(td, "BadF.ChF")
        open(p, "w").write(bad_src)
        good, bad = chf2py.translate_file(p, 2)
    assert "testkernel" in bad and "testkernel" not in good and "sections" in good
