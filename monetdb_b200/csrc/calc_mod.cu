// calc_mod.cu — instantiates the BATcalcmod type matrix (gdk_calc_mod.c:121-290; see
// calc.cuh); one translation unit per operator keeps the nvcc jobs parallel.
#include "calc.cuh"
int calc_mod_entry(int t1, int t2, int t3, const CalcOperand &o1, const CalcOperand &o2, void *dst, uint64_t n, CalcOut &out)
{
	return calc_dispatch<OP_MOD>(t1, t2, t3, o1, o2, dst, n, out);
}
