// calc.cu — host drivers of BATcalcadd/sub/mul/div/mod and their constant-operand
// forms (gdk_calc_addsub.c:1480-1664,3166-3330; gdk_calc_mul.c:2020-2210;
// gdk_calc_div.c:1945-2060; gdk_calc_mod.c:1179-1290): candidate set-up, result allocation, error text and
// result properties.  The kernels live in calc.cuh / calc_<op>.cu.
#include "calc.cuh"

static const char *opsym[] = {"+", "-", "*", "/", "%"};
static const char *opname[] = {"add", "sub", "mul", "div", "mod"};
static calc_fn opentry[] = {calc_add_entry, calc_sub_entry, calc_mul_entry, calc_div_entry, calc_mod_entry};

// sign of a constant (VARcalcsign) for the mul/div sortedness rules
static int cst_sign(int t, const void *v)
{
	switch (t) {
	case B200_bte: return (*(const bte *) v > 0) - (*(const bte *) v < 0);
	case B200_sht: return (*(const sht *) v > 0) - (*(const sht *) v < 0);
	case B200_int: return (*(const int *) v > 0) - (*(const int *) v < 0);
	case B200_lng: return (*(const lng *) v > 0) - (*(const lng *) v < 0);
	case B200_hge: return (*(const hge *) v > 0) - (*(const hge *) v < 0);
	case B200_flt: return (*(const flt *) v > 0) - (*(const flt *) v < 0);
	case B200_dbl: return (*(const dbl *) v > 0) - (*(const dbl *) v < 0);
	}
	return 0;
}

// mode: 0 = b1 OP b2, 1 = b1 OP cst, 2 = cst OP b1
static b200_bat *calc_binary(int op, int mode, const b200_bat *b1, const b200_bat *b2, const b200_bat *s1,
			     const b200_bat *s2, const void *cst, int csttype, int tp, const char *func)
{
	if (b1 == nullptr || (mode == 0 && b2 == nullptr) || (mode != 0 && cst == nullptr)) {
		set_err("%s: BAT required\n", func);
		return nullptr;
	}
	Staged sb1, sb2, ss1, ss2;
	if (!stage(b1, sb1) || !stage(b2, sb2) || !stage(s1, ss1) || !stage(s2, ss2))
		return nullptr;
	b1 = sb1.b;
	b2 = sb2.b;
	Cand ci1, ci2;
	if (!cand_init(ci1, b1, ss1.b))
		return nullptr;
	if (mode == 0) {
		if (!cand_init(ci2, b2, ss2.b))
			return nullptr;
		if (ci1.ncand != ci2.ncand || ci1.rhseq != ci2.rhseq) {
			/* gdk_calc_addsub.c:1494, gdk_calc_mul.c:2041 */
			if (op >= OP_MUL)
				set_err("%s: inputs not the same size.\n", func);
			else
				set_err("inputs not the same size.\n");
			return nullptr;
		}
	}
	const int st3 = basetype(tp);
	if (atomsize(st3) <= 0 || st3 == B200_void) {
		set_err("%s: type combination (%s(%s,%s)->%s) not supported.\n", func, opname[op],
			atomname(b1->ttype), atomname(mode == 0 ? b2->ttype : csttype), atomname(tp));
		return nullptr;
	}
	b200_bat *bn = bat_new(ci1.rhseq, tp, ci1.ncand);
	if (bn == nullptr)
		return nullptr;
	if (ci1.ncand == 0)
		return bn;
	if (b1->ttype == B200_void || (b2 && b2->ttype == B200_void)) {
		b200_BATfree(bn);
		set_err("%s: void columns are not on the device path\n", func);
		return nullptr;
	}
	CalcOperand o1, o2;
	const int tb1 = basetype(b1->ttype);
	if (mode == 0) {
		o1 = CalcOperand{b1, &ci1, nullptr, tb1};
		o2 = CalcOperand{b2, &ci2, nullptr, basetype(b2->ttype)};
	} else if (mode == 1) {
		o1 = CalcOperand{b1, &ci1, nullptr, tb1};
		o2 = CalcOperand{nullptr, nullptr, cst, basetype(csttype)};
	} else {
		o1 = CalcOperand{nullptr, nullptr, cst, basetype(csttype)};
		o2 = CalcOperand{b1, &ci1, nullptr, tb1};
	}
	CalcOut out;
	ctx()->algo = "calc: cuda streaming";
	int rc = opentry[op](o1.type, o2.type, st3, o1, o2, bn->base, ci1.ncand, out);
	if (rc <= 0) {
		b200_BATfree(bn);
		if (rc < 0)
			set_err("%s: type combination (%s(%s,%s)->%s) not supported.\n", func, opname[op],
				atomname(o1.type), atomname(o2.type), atomname(tp));
		return nullptr;
	}
	if (out.fail != ~0ULL) {
		b200_BATfree(bn);
		switch (out.status) {
		case ST_DIVZERO:
			set_err("22012!division by zero.\n");
			break;
		case ST_DIVRANGE:
			/* the reference returns NULL here without a message (BUN_NONE + 2) */
			set_err("22003!overflow in calculation (division result out of range).\n");
			break;
		default:
			set_err("22003!overflow in calculation %s%s%s.\n", out.l, opsym[op], out.r);
		}
		return nullptr;
	}
	const uint64_t n = ci1.ncand, nils = out.nils;
	bn->count = n;
	bn->tnil = nils != 0;
	bn->tnonil = nils == 0;
	bn->tkey = n <= 1;
	const bool triv = n <= 1 || nils == n;
	bool srt = false, rev = false;
	const bool s1s = b1->tsorted, s1r = b1->trevsorted;
	/* sortedness rules of the reference drivers; an oid-list candidate keeps order */
	if (op == OP_ADD) {
		if (mode == 0) {
			srt = s1s && b2->tsorted && nils == 0;
			rev = s1r && b2->trevsorted && nils == 0;
		} else {
			srt = s1s && nils == 0;
			rev = s1r && nils == 0;
		}
	} else if (op == OP_SUB) {
		if (mode == 1) {
			srt = s1s && nils == 0;
			rev = s1r && nils == 0;
		} else if (mode == 2) {
			srt = s1r && nils == 0;
			rev = s1s && nils == 0;
		}
	} else if (op == OP_MUL && mode != 0) {
		int sg = cst_sign(basetype(csttype), cst);
		srt = (sg >= 0 && s1s && nils == 0) || (sg <= 0 && s1r && nils == 0);
		rev = (sg >= 0 && s1r && nils == 0) || (sg <= 0 && s1s && nils == 0);
	} else if (op == OP_DIV && mode == 1) {
		int sg = cst_sign(basetype(csttype), cst);
		srt = (sg > 0 && s1s && nils == 0) || (sg < 0 && s1r && nils == 0);
		rev = (sg > 0 && s1r && nils == 0) || (sg < 0 && s1s && nils == 0);
	}
	bn->tsorted = srt || triv;
	bn->trevsorted = rev || triv;
	return bn;
}

#define BINOP(NAME, OP)                                                                                    \
	extern "C" b200_bat *b200_BATcalc##NAME(const b200_bat *b1, const b200_bat *b2, const b200_bat *s1, \
						const b200_bat *s2, int tp)                                 \
	{                                                                                                  \
		b200_clrerr();                                                                             \
		return calc_binary(OP, 0, b1, b2, s1, s2, nullptr, 0, tp, "BATcalc" #NAME);                \
	}                                                                                                  \
	extern "C" b200_bat *b200_BATcalc##NAME##cst(const b200_bat *b, const void *v, int vtype,          \
						     const b200_bat *s, int tp)                             \
	{                                                                                                  \
		b200_clrerr();                                                                             \
		return calc_binary(OP, 1, b, nullptr, s, nullptr, v, vtype, tp, "BATcalc" #NAME "cst");    \
	}                                                                                                  \
	extern "C" b200_bat *b200_BATcalccst##NAME(const void *v, int vtype, const b200_bat *b,            \
						   const b200_bat *s, int tp)                               \
	{                                                                                                  \
		b200_clrerr();                                                                             \
		return calc_binary(OP, 2, b, nullptr, s, nullptr, v, vtype, tp, "BATcalccst" #NAME);       \
	}
BINOP(add, OP_ADD)
BINOP(sub, OP_SUB)
BINOP(mul, OP_MUL)
BINOP(div, OP_DIV)
BINOP(mod, OP_MOD)
