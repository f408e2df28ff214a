// Tail of path_kernel: control points out, power-basis coefficients, arc-length tables (A4),
// path length / status, and the curve point + curvature at every toppra gridpoint (A16/A17).
// Included inside the kernel; see path_kernel.cu for the names in scope.
  const int nc = nseg * NC, nt = nseg * kTable;
  // ---- control points out: the CTA's slice of [P][nc] is contiguous
  for (int it = tid; it < PBa * nc; it += T) {
    const bool b = bad[it / nc] != 0;
    if (g.ctrl_x) ((R*)g.ctrl_x)[(size_t)p0 * nc + it] = b ? (R)0 : cx[it];
    if (g.ctrl_y) ((R*)g.ctrl_y)[(size_t)p0 * nc + it] = b ? (R)0 : cy[it];
  }
  // ---- power basis per segment: a_k = C(d,k) Delta^k P_0 (curve), h_k likewise for the hodograph
  //      control points d (P[j+1] - P[j])  (sc:1041-1053)
  for (int it = tid; it < PBa * nseg; it += T) {
    R dx[NC], dy[NC];
#pragma unroll
    for (int j = 0; j < NC; ++j) { dx[j] = cx[NC * it + j]; dy[j] = cy[NC * it + j]; }
    R hdx[DEG], hdy[DEG];
#pragma unroll
    for (int j = 0; j < DEG; ++j) { hdx[j] = (R)DEG * (dx[j + 1] - dx[j]); hdy[j] = (R)DEG * (dy[j + 1] - dy[j]); }
    R binom = 1;
#pragma unroll
    for (int k = 0; k <= DEG; ++k) {
      ax[NC * it + k] = binom * dx[0]; ay[NC * it + k] = binom * dy[0];
#pragma unroll
      for (int j = 0; j + k < DEG; ++j) { dx[j] = dx[j + 1] - dx[j]; dy[j] = dy[j + 1] - dy[j]; }
      binom = binom * (R)(DEG - k) / (R)(k + 1);
    }
    binom = 1;
#pragma unroll
    for (int k = 0; k < DEG; ++k) {
      hx[DEG * it + k] = binom * hdx[0]; hy[DEG * it + k] = binom * hdy[0];
#pragma unroll
      for (int j = 0; j + k < DEG - 1; ++j) { hdx[j] = hdx[j + 1] - hdx[j]; hdy[j] = hdy[j + 1] - hdy[j]; }
      binom = binom * (R)(DEG - 1 - k) / (R)(k + 1);
    }
  }
  __syncthreads();

  // ---- A4: arc-length pieces, one item per (path, segment, sub-interval): 32 nodes in table order
  //      (sc:854-880).  Pieces go to tab[..][k + 1]; the cumulative sum follows.
  R* tab = g.table_in_smem ? sm.take((size_t)PB * nt) : (R*)g.table + (size_t)p0 * nt;
  for (int it = tid; it < PBa * nseg * kSub; it += T) {
    const int sg = it / kSub, k = it - sg * kSub;
    const R a = fmin((R)k * (R)0.01, (R)1), b = fmin((R)(k + 1) * (R)0.01, (R)1);
    const R half = (b - a) / 2, mid = (b + a) / 2;
    R kx[DEG], ky[DEG];
#pragma unroll
    for (int j = 0; j < DEG; ++j) { kx[j] = hx[DEG * sg + j]; ky[j] = hy[DEG * sg + j]; }
    R acc = 0;
#pragma unroll 8
    for (int i = 0; i < 32; ++i) {
      const R u = half * (R)GL32_X[i] + mid;
      R dx = kx[DEG - 1], dy = ky[DEG - 1];
#pragma unroll
      for (int j = DEG - 2; j >= 0; --j) { dx = dx * u + kx[j]; dy = dy * u + ky[j]; }
      const R n2 = dx * dx + dy * dy;
      acc += (R)GL32_W[i] * (n2 > 0 ? n2 * rsqrt_<R>(n2) : (R)0);
    }
    tab[sg * kTable + k + 1] = acc * half;
  }
  __syncthreads();
  // cumulative table per segment, in order (sc:882-886)
  for (int it = tid; it < PBa * nseg; it += T) {
    R* c = tab + it * kTable;
    R sum = 0;
    c[0] = 0;
    for (int k = 1; k <= kSub; ++k) { sum += c[k]; c[k] = sum; }
    segtot[it] = sum;
  }
  __syncthreads();
  if (tid < PBa) {
    const int p = p0 + tid;
    double acc = 0;
    for (int s = 0; s < nseg; ++s) { acc += (double)segtot[tid * nseg + s]; segend[tid * nseg + s] = acc; }
    const bool ok = !bad[tid] && acc > 0 && isfinite(acc);
    if (!ok) bad[tid] = 1;
    scal[tid * 8 + 4] = (R)acc;
    scal[tid * 8 + 5] = g.vel_start ? ((const R*)g.vel_start)[p] : (R)0;
    scal[tid * 8 + 6] = g.vel_end ? ((const R*)g.vel_end)[p] : (R)0;
    ((R*)g.L)[p] = ok ? (R)acc : (R)0;
    if (g.cost) ((R*)g.cost)[p] = ok ? scal[tid * 8 + 1] : (R)0;
    g.status[p] = ok ? SCB_ST_OK : SCB_ST_ERR_BAD_PATH;
  }
  __syncthreads();
  if (g.table && g.table_in_smem) {
    R* o_tab = (R*)g.table + (size_t)p0 * nt;
    for (int it = tid; it < PBa * nt; it += T) o_tab[it] = bad[it / nt] ? (R)0 : tab[it];
  } else if (g.table) {
    R* o_tab = (R*)g.table + (size_t)p0 * nt;
    for (int it = tid; it < PBa * nt; it += T) if (bad[it / nt]) o_tab[it] = 0;
  }

  // ---- geometry at the toppra gridpoints: q(s_i) -> (segment, u) -> point, curvature.
  //      Items are (gridpoint, path) with the path index fastest, so a warp's stores to the
  //      [G][P] arrays are PB-contiguous.
  if (g.px || g.py || g.kappa) {
    const int G = g.G, N = G - 1;
    const double step = 1.0 / (double)N;
    R* o_px = (R*)g.px; R* o_py = (R*)g.py; R* o_k = (R*)g.kappa;
    for (int it = tid; it < PBa * G; it += T) {
      const int i = it / PBa, pl = it - i * PBa;
      const size_t o = (size_t)i * P + p0 + pl;
      if (bad[pl]) {
        if (o_px) o_px[o] = 0; if (o_py) o_py[o] = 0; if (o_k) o_k[o] = 0;
        continue;
      }
      // q(s_i) in FP64 against the FP64 segment ends, whatever R is (three DFMAs per gridpoint)
      HermiteQ<double> hq;
      const double* se_ = segend + pl * nseg;
      const double Ltot = se_[nseg - 1];
      hq.init(Ltot, (double)scal[pl * 8 + 5], (double)scal[pl * 8 + 6]);
      double q = hq.q(i == N ? 1.0 : (double)i * step);
      q = fmin(fmax(q, 0.0), Ltot);
      // segment: first j with q <= segend[j]
      int lo = 0, hi = nseg - 1;
      while (lo < hi) { const int m = (lo + hi) >> 1; if (q <= se_[m]) hi = m; else lo = m + 1; }
      const int sgl = lo, sg = pl * nseg + sgl;
      const R r = (R)(q - (sgl > 0 ? se_[sgl - 1] : 0.0));
      const R* c = tab + sg * kTable;
      int kl = 0, kh = kSub;
      while (kh - kl > 1) { const int m = (kl + kh) >> 1; if (c[m] <= r) kl = m; else kh = m; }
      const R h = (R)0.01, a0 = fmin((R)kl * h, (R)1);
      const R ck = c[kl], target = r - ck;
      R du = h * target / (c[kl + 1] - ck);
      du = fmin(fmax(du, (R)0), h);
      R kx[DEG], ky[DEG];  // B' in power basis: (k+1) a_{k+1}
#pragma unroll
      for (int j = 0; j < DEG; ++j) { kx[j] = (R)(j + 1) * ax[NC * sg + j + 1]; ky[j] = (R)(j + 1) * ay[NC * sg + j + 1]; }
      auto speed = [&](R u) {
        R dx = kx[DEG - 1], dy = ky[DEG - 1];
#pragma unroll
        for (int j = DEG - 2; j >= 0; --j) { dx = dx * u + kx[j]; dy = dy * u + ky[j]; }
        const R n2 = dx * dx + dy * dy;
        return n2 > 0 ? n2 * rsqrt_<R>(n2) : (R)0;
      };
#pragma unroll 1
      for (int nit = 0; nit < kInvNewton; ++nit) {
        const R half = (R)0.5 * du, mid = a0 + half;
        R F = 0;
#pragma unroll
        for (int gq = 0; gq < 4; ++gq) F += (R)kPGL4W[gq] * speed(mid + half * (R)kPGL4X[gq]);
        F *= half;
        du = du - (F - target) / speed(a0 + du);
        du = fmin(fmax(du, (R)0), h);
      }
      const R u = fmin(a0 + du, (R)1);
      R bx = ax[NC * sg + DEG], by = ay[NC * sg + DEG];
#pragma unroll
      for (int j = DEG - 1; j >= 0; --j) { bx = bx * u + ax[NC * sg + j]; by = by * u + ay[NC * sg + j]; }
      R x1 = kx[DEG - 1], y1 = ky[DEG - 1];
#pragma unroll
      for (int j = DEG - 2; j >= 0; --j) { x1 = x1 * u + kx[j]; y1 = y1 * u + ky[j]; }
      R x2 = (R)(DEG - 1) * kx[DEG - 1], y2 = (R)(DEG - 1) * ky[DEG - 1];
#pragma unroll
      for (int j = DEG - 3; j >= 0; --j) { x2 = x2 * u + (R)(j + 1) * kx[j + 1]; y2 = y2 * u + (R)(j + 1) * ky[j + 1]; }
      const R n2 = x1 * x1 + y1 * y1, rn = rsqrt_<R>(n2);
      if (o_px) o_px[o] = bx;
      if (o_py) o_py[o] = by;
      if (o_k) o_k[o] = (x1 * y2 - y1 * x2) * (rn * rn * rn);  // sc:1029
    }
  }
