// aggr_init.cuh — BATgroupaggrinit (gdk_aggr.c:65-146) shared by the grouped aggregates
#pragma once
#include "common.cuh"

// ---------------------------------------------------------------- BATgroupaggrinit (aggr.cu)
struct AggrSetup {
	Staged sb, sg, se, ss;
	Cand ci;
	const b200_bat *b = nullptr, *g = nullptr, *e = nullptr;
	oid min = 0, max = 0;
	uint64_t ngrp = 0;
};

bool aggr_init(AggrSetup &a, const b200_bat *b, const b200_bat *g, const b200_bat *e, const b200_bat *s);
