// Body of path_kernel<R, 5>: Lau09 quintic builder (A5), curvature/arc-length cost with analytic
// gradient (A6) and the iRPROP- loop (A7).  Included inside the kernel; see path_kernel.cu for the
// names in scope.  tx/ty hold the frozen base tangents Bt.  No reference code exists for these
// stages: the frozen definition is DESIGN.md "New stages" (DESIGN.md "New stages").
    R* w0x = sm.take(PB * W); R* w0y = sm.take(PB * W); R* mag = sm.take(PB * W); R* alpha = sm.take(PB * W);
    R* e = sm.take(PB * W); R* se = sm.take(PB * W); R* pe = sm.take(PB * W);
    R* swx = sm.take(PB * W); R* swy = sm.take(PB * W); R* pwx = sm.take(PB * W); R* pwy = sm.take(PB * W);
    R* be = sm.take(PB * W); R* bwx = sm.take(PB * W); R* bwy = sm.take(PB * W);
    R* Tx = sm.take(PB * W); R* Ty = sm.take(PB * W); R* Ax = sm.take(PB * W); R* Ay = sm.take(PB * W);
    R* gTx = sm.take(PB * W); R* gTy = sm.take(PB * W); R* gAx = sm.take(PB * W); R* gAy = sm.take(PB * W);
    R* gWx = sm.take(PB * W); R* gWy = sm.take(PB * W); R* ge = sm.take(PB * W);
    R* gx = sm.take(PB * nseg * 6); R* gy = sm.take(PB * nseg * 6); R* Js = sm.take(PB * nseg);
    R* nb = sm.take(32 * 13);  // per Gauss-Legendre node on [0,1]: c1[6], c2[6], wq
    const scb_opt& o = g.opt;
    for (int it = tid; it < PBa * W; it += T) {
      const int pl = it / W, i = it - pl * W;
      w0x[it] = wx[it]; w0y[it] = wy[it];
      mag[it] = sqrt(tx[it] * tx[it] + ty[it] * ty[it]);
      R al = 0;
      if (i > 0 && i < W - 1) {
        const R dp = dist2<R>(wx[it - 1], wy[it - 1], wx[it], wy[it]), dn = dist2<R>(wx[it], wy[it], wx[it + 1], wy[it + 1]);
        al = dn / (dp + dn);
      }
      alpha[it] = al;
      e[it] = (R)o.e_init; se[it] = (R)o.step_e; pe[it] = 0;
      swx[it] = swy[it] = (R)o.step_w * mag[it]; pwx[it] = pwy[it] = 0;
    }
    if (tid < 32) {  // basis values of B', B'' at node `tid`
      const double t = 0.5 * GL32_X[tid] + 0.5, s = 1 - t;
      const double b4[5] = {s * s * s * s, 4 * t * s * s * s, 6 * t * t * s * s, 4 * t * t * t * s, t * t * t * t};
      const double b3[4] = {s * s * s, 3 * t * s * s, 3 * t * t * s, t * t * t};
#pragma unroll
      for (int j = 0; j < 6; ++j) {
        const double lo = j > 0 ? b4[j - 1] : 0, hi = j < 5 ? b4[j] : 0;
        nb[tid * 13 + j] = (R)(5 * (lo - hi));
        const double a = j >= 2 ? b3[j - 2] : 0, b = (j >= 1 && j <= 4) ? b3[j - 1] : 0, c = j <= 3 ? b3[j] : 0;
        nb[tid * 13 + 6 + j] = (R)(20 * (a - 2 * b + c));
      }
      nb[tid * 13 + 12] = (R)(0.5 * GL32_W[tid]);
    }
    __syncthreads();
    const R lambda = (R)o.lambda;
    auto build = [&]() {
      for (int it = tid; it < PBa * W; it += T) { Tx[it] = e[it] * tx[it]; Ty[it] = e[it] * ty[it]; }
      __syncthreads();
      for (int it = tid; it < PBa * W; it += T) {
        const int pl = it / W, i = it - pl * W;
        R px = 0, py = 0, nx = 0, ny = 0;
        if (i > 0) {
          px = (R)6 * (wx[it - 1] - wx[it]) + (R)4 * Tx[it] + (R)2 * Tx[it - 1];
          py = (R)6 * (wy[it - 1] - wy[it]) + (R)4 * Ty[it] + (R)2 * Ty[it - 1];
        }
        if (i + 1 < W) {
          nx = (R)6 * (wx[it + 1] - wx[it]) - (R)4 * Tx[it] - (R)2 * Tx[it + 1];
          ny = (R)6 * (wy[it + 1] - wy[it]) - (R)4 * Ty[it] - (R)2 * Ty[it + 1];
        }
        if (i == 0) { Ax[it] = nx; Ay[it] = ny; }
        else if (i + 1 == W) { Ax[it] = px; Ay[it] = py; }
        else { Ax[it] = alpha[it] * px + ((R)1 - alpha[it]) * nx; Ay[it] = alpha[it] * py + ((R)1 - alpha[it]) * ny; }
      }
      __syncthreads();
      for (int it = tid; it < PBa * nseg; it += T) {
        const int pl = it / nseg, s = it - pl * nseg, i = pl * W + s;
        const R fifth = (R)1 / (R)5, twentieth = (R)1 / (R)20;
        const R p0x = wx[i], p0y = wy[i], p5x = wx[i + 1], p5y = wy[i + 1];
        const R p1x = p0x + fifth * Tx[i], p1y = p0y + fifth * Ty[i];
        const R p4x = p5x - fifth * Tx[i + 1], p4y = p5y - fifth * Ty[i + 1];
        R* ox = cx + 6 * it; R* oy = cy + 6 * it;
        ox[0] = p0x; oy[0] = p0y; ox[1] = p1x; oy[1] = p1y;
        ox[2] = twentieth * Ax[i] + (R)2 * p1x - p0x; oy[2] = twentieth * Ay[i] + (R)2 * p1y - p0y;
        ox[3] = twentieth * Ax[i + 1] + (R)2 * p4x - p5x; oy[3] = twentieth * Ay[i + 1] + (R)2 * p4y - p5y;
        ox[4] = p4x; oy[4] = p4y; ox[5] = p5x; oy[5] = p5y;
      }
      __syncthreads();
    };
    const int K = o.iters;
    const int n_cg = ((PBa * nseg * kNG + 31) / 32) * 32;  // whole warps so the butterfly below is convergent
    for (int iter = 0; iter <= K; ++iter) {
      build();
      // ---- A6: cost and control-point gradient; kNG lanes per segment, kNodesPerLane nodes each
      for (int it = tid; it < n_cg; it += T) {
        const bool live = it < PBa * nseg * kNG;
        const int sg = live ? it / kNG : 0, grp = it % kNG;
        R Px[6], Py[6];
#pragma unroll
        for (int j = 0; j < 6; ++j) { Px[j] = cx[6 * sg + j]; Py[j] = cy[6 * sg + j]; }
        R J = 0, ggx[6], ggy[6];
#pragma unroll
        for (int j = 0; j < 6; ++j) { ggx[j] = 0; ggy[j] = 0; }
        const bool want_grad = iter < K;
#pragma unroll 2
        for (int qn = 0; qn < kNodesPerLane; ++qn) {
          const R* b = nb + (grp * kNodesPerLane + qn) * 13;
          R x1 = 0, y1 = 0, x2 = 0, y2 = 0;
#pragma unroll
          for (int j = 0; j < 6; ++j) {
            x1 += b[j] * Px[j]; y1 += b[j] * Py[j]; x2 += b[6 + j] * Px[j]; y2 += b[6 + j] * Py[j];
          }
          // n = |B'|; every power of 1/n comes from one rsqrt (no divisions in the node loop)
          const R n2 = x1 * x1 + y1 * y1, rn = rsqrt_<R>(n2), n = n2 * rn, rn2 = rn * rn;
          const R c = x1 * y2 - y1 * x2;
          const R in5 = (rn2 * rn2) * rn;
          const R wq = b[12];
          J += wq * (c * c * in5 + lambda * n);
          if (want_grad) {
            const R k1 = (R)2 * c * in5, k2 = ((R)5 * c * c * in5) * rn2, k3 = lambda * rn;
            const R fx1 = wq * (k1 * y2 - k2 * x1 + k3 * x1), fy1 = wq * (-k1 * x2 - k2 * y1 + k3 * y1);
            const R fx2 = wq * (-k1 * y1), fy2 = wq * (k1 * x1);
#pragma unroll
            for (int j = 0; j < 6; ++j) { ggx[j] += fx1 * b[j] + fx2 * b[6 + j]; ggy[j] += fy1 * b[j] + fy2 * b[6 + j]; }
          }
        }
#pragma unroll
        for (int off = 1; off < kNG; off <<= 1) {
          J += __shfl_xor_sync(kFull, J, off);
          if (want_grad) {
#pragma unroll
            for (int j = 0; j < 6; ++j) { ggx[j] += __shfl_xor_sync(kFull, ggx[j], off); ggy[j] += __shfl_xor_sync(kFull, ggy[j], off); }
          }
        }
        if (live && grp == 0) {
          Js[sg] = J;
          if (want_grad) {
#pragma unroll
            for (int j = 0; j < 6; ++j) { gx[6 * sg + j] = ggx[j]; gy[6 * sg + j] = ggy[j]; }
          }
        }
      }
      __syncthreads();
      if (tid < PBa) {  // J of each path: segments in order; best-so-far bookkeeping
        R J = 0;
        for (int s = 0; s < nseg; ++s) J += Js[tid * nseg + s];
        const bool better = iter == 0 || J < scal[tid * 8 + 1];
        if (better) scal[tid * 8 + 1] = J;
        improve[tid] = better ? 1 : 0;
      }
      __syncthreads();
      // ---- chain rule to (T, A, W) per waypoint, gathering from the two adjacent segments
      for (int it = tid; it < PBa * W; it += T) {
        const int pl = it / W, i = it - pl * W;
        if (improve[pl]) { be[it] = e[it]; bwx[it] = wx[it]; bwy[it] = wy[it]; }
        if (iter == K) continue;
        R wxs = 0, wys = 0, txs = 0, tys = 0, axs = 0, ays = 0;
        if (i > 0) {
          const R* sx = gx + 6 * (pl * nseg + i - 1); const R* sy = gy + 6 * (pl * nseg + i - 1);
          wxs += (sx[3] + sx[4] + sx[5]); wys += (sy[3] + sy[4] + sy[5]);
          txs -= ((R)0.4 * sx[3] + (R)0.2 * sx[4]); tys -= ((R)0.4 * sy[3] + (R)0.2 * sy[4]);
          axs += (R)0.05 * sx[3]; ays += (R)0.05 * sy[3];
        }
        if (i < nseg) {
          const R* sx = gx + 6 * (pl * nseg + i); const R* sy = gy + 6 * (pl * nseg + i);
          wxs += (sx[0] + sx[1] + sx[2]); wys += (sy[0] + sy[1] + sy[2]);
          txs += ((R)0.2 * sx[1] + (R)0.4 * sx[2]); tys += ((R)0.2 * sy[1] + (R)0.4 * sy[2]);
          axs += (R)0.05 * sx[2]; ays += (R)0.05 * sy[2];
        }
        gWx[it] = wxs; gWy[it] = wys; gTx[it] = txs; gTy[it] = tys; gAx[it] = axs; gAy[it] = ays;
      }
      if (iter == K) break;
      __syncthreads();
      for (int it = tid; it < PBa * W; it += T) {
        const int pl = it / W, i = it - pl * W;
        auto ap = [&](int k) { return k == 0 ? (R)0 : (k == W - 1 ? (R)1 : alpha[pl * W + k]); };
        auto an = [&](int k) { return k == W - 1 ? (R)0 : (k == 0 ? (R)1 : (R)1 - alpha[pl * W + k]); };
        R wxs = gWx[it], wys = gWy[it], txs = gTx[it], tys = gTy[it];
        if (i > 0) {  // waypoint i is the "next" of i-1
          const R k6 = (R)6 * an(i - 1), k2 = (R)2 * an(i - 1);
          wxs += k6 * gAx[it - 1]; wys += k6 * gAy[it - 1]; txs -= k2 * gAx[it - 1]; tys -= k2 * gAy[it - 1];
        }
        {
          const R k6 = (R)6 * (ap(i) + an(i)), k4 = (R)4 * (ap(i) - an(i));
          wxs -= k6 * gAx[it]; wys -= k6 * gAy[it]; txs += k4 * gAx[it]; tys += k4 * gAy[it];
        }
        if (i + 1 < W) {  // waypoint i is the "prev" of i+1
          const R k6 = (R)6 * ap(i + 1), k2 = (R)2 * ap(i + 1);
          wxs += k6 * gAx[it + 1]; wys += k6 * gAy[it + 1]; txs += k2 * gAx[it + 1]; tys += k2 * gAy[it + 1];
        }
        ge[it] = txs * tx[it] + tys * ty[it];
        Tx[it] = wxs; Ty[it] = wys;  // final dJ/dW; Tx/Ty are free until the next build()
      }
      __syncthreads();
      if (tid < PBa) {
        R gemax = 0, gwmax = 0;
        for (int i = 0; i < W; ++i) {
          gemax = fmax(gemax, fabs(ge[tid * W + i]));
          if (i > 0 && i + 1 < W) gwmax = fmax(gwmax, fmax(fabs(Tx[tid * W + i]), fabs(Ty[tid * W + i])));
        }
        scal[tid * 8 + 2] = gemax; scal[tid * 8 + 3] = gwmax;
      }
      __syncthreads();
      // ---- A7: iRPROP-
      auto rprop1 = [](R gv, R& gprev, R& step, R smin, R smax) -> R {
        const R s = gv * gprev;
        if (s > 0) step = fmin(step * (R)1.2, smax);
        else if (s < 0) { step = fmax(step * (R)0.5, smin); gv = 0; }
        gprev = gv;
        return gv > 0 ? -step : (gv < 0 ? step : (R)0);
      };
      for (int it = tid; it < PBa * W; it += T) {
        const int pl = it / W, i = it - pl * W;
        const R te = (R)1e-9 * scal[pl * 8 + 2], tw = (R)1e-9 * scal[pl * 8 + 3];
        const R gi = fabs(ge[it]) <= te ? (R)0 : ge[it];
        const R ei = e[it] + rprop1(gi, pe[it], se[it], (R)1e-6, (R)1.0);
        e[it] = fmin((R)o.e_max, fmax((R)o.e_min, ei));
        if (o.move_waypoints && i > 0 && i + 1 < W) {
          const R gvx = fabs(Tx[it]) <= tw ? (R)0 : Tx[it], gvy = fabs(Ty[it]) <= tw ? (R)0 : Ty[it];
          const R smin = (R)1e-6 * mag[it], smax = (R)0.5 * mag[it], r = (R)o.w_radius;
          const R nx = wx[it] + rprop1(gvx, pwx[it], swx[it], smin, smax);
          const R ny = wy[it] + rprop1(gvy, pwy[it], swy[it], smin, smax);
          wx[it] = fmin(w0x[it] + r, fmax(w0x[it] - r, nx));
          wy[it] = fmin(w0y[it] + r, fmax(w0y[it] - r, ny));
        }
      }
      __syncthreads();
    }
    __syncthreads();
    for (int it = tid; it < PBa * W; it += T) { e[it] = be[it]; wx[it] = bwx[it]; wy[it] = bwy[it]; }
    __syncthreads();
    build();
