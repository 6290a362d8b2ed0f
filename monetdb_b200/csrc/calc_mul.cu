// calc_mul.cu — instantiates the BATcalcmul type matrix (see calc.cuh); one
// translation unit per operator keeps the nvcc jobs parallel.
#include "calc.cuh"
int calc_mul_entry(int t1, int t2, int t3, const CalcOperand &o1, const CalcOperand &o2, void *dst, uint64_t n, CalcOut &out)
{
	return calc_dispatch<OP_MUL>(t1, t2, t3, o1, o2, dst, n, out);
}
