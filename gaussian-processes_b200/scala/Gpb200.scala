// Gpb200.scala -- reference-side binding of libgpb200.so (include/gpb200.h) through the Java
// Foreign Function & Memory API (Panama, JDK 22+; no C glue).  SOURCE ONLY: this image has no
// JVM/scalac, so this file is not compiled or tested here (DESIGN.md section 1).  It shows how the
// reference's public methods keep their signatures while the hot path runs on the B200.
//
// Drop-in points (core/src/main/scala/gp/):
//   GaussianProcess.loglikelihood  GaussianProcess.scala:109-127  -> Gpb200.loglikelihood
//   Predict.fit                    Predict.scala:57-94            -> Gpb200.fit
//   KernelParameters.sample        KernelParameters.scala:231-246 -> unchanged (it calls loglikelihood)
//   MeanParameters.samplePlane     MeanParameters.scala:74-103    -> Gpb200.samplePlane
//   GaussianProcess.draw           GaussianProcess.scala:67-83    -> Gpb200.draw
//   GaussianProcess.mllGradient    GaussianProcess.scala:166-187  -> Gpb200.mllGradient
//   Mcmc.sample (many chains)      Mcmc.scala:63-76               -> Gpb200.mhRunEx (device resident)
//   Streaming.writeParallelChain   SimulatedGp.scala:163-165      -> Gpb200.commUniqueId / commInitRank /
//                                                                    commShard + mhRunEx(gather = 1)
package com.github.jonnylaw.gp

import java.lang.foreign._
import java.lang.foreign.ValueLayout._
import java.lang.invoke.MethodHandle
import breeze.stats.distributions.Gaussian
import GaussianProcess.{Data, Parameters}

object Gpb200 {
  private val linker = Linker.nativeLinker()
  private val lib = SymbolLookup.libraryLookup("libgpb200.so", Arena.global())
  private def fn(name: String, fd: FunctionDescriptor): MethodHandle =
    linker.downcallHandle(lib.find(name).orElseThrow(), fd)

  // struct gpb_term { int kind; double p0, p1, p2; }   (32 bytes, 8-aligned)
  private val TERM = MemoryLayout.structLayout(JAVA_INT.withName("kind"), MemoryLayout.paddingLayout(4),
    JAVA_DOUBLE.withName("p0"), JAVA_DOUBLE.withName("p1"), JAVA_DOUBLE.withName("p2"))
  // struct gpb_ll { double Q, U1, S; int n; int info; } (32 bytes)
  private val LL = MemoryLayout.structLayout(JAVA_DOUBLE.withName("Q"), JAVA_DOUBLE.withName("U1"),
    JAVA_DOUBLE.withName("S"), JAVA_INT.withName("n"), JAVA_INT.withName("info"))

  private val ctxCreate = fn("gpb_ctx_create", FunctionDescriptor.of(JAVA_INT, JAVA_INT, ADDRESS))
  private val dsCreate = fn("gpb_dataset_create",
    FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT, JAVA_INT, ADDRESS, ADDRESS, ADDRESS))
  private val dsDestroy = fn("gpb_dataset_destroy", FunctionDescriptor.ofVoid(ADDRESS))
  private val loglik = fn("gpb_loglik", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS,
    JAVA_INT, ADDRESS, JAVA_INT, JAVA_DOUBLE, ADDRESS))
  // gpb_loglik_noalloc(device, stream, n, dim, d_x, d_y, terms, nterms, beta, nbeta, nugget, out): the same
  // evaluation for callers that may not allocate (static device scratch, n <= 512); d_x / d_y are DEVICE
  // addresses the caller owns (e.g. from its own CUDA binding) -- no counterpart in the reference
  private val loglikNoAlloc = fn("gpb_loglik_noalloc", FunctionDescriptor.of(JAVA_INT, JAVA_INT, ADDRESS, JAVA_INT,
    JAVA_INT, ADDRESS, ADDRESS, ADDRESS, JAVA_INT, ADDRESS, JAVA_INT, JAVA_DOUBLE, ADDRESS))
  private val llValue = fn("gpb_ll_value", FunctionDescriptor.of(JAVA_DOUBLE, ADDRESS, JAVA_INT))
  private val predict = fn("gpb_predict", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS,
    JAVA_INT, JAVA_DOUBLE, JAVA_INT, ADDRESS, ADDRESS, ADDRESS, ADDRESS))

  // one context per JVM thread (a gpb_ctx is not shared between threads: SimulatedGp.scala:163-165
  // runs 2 chains on 2 threads)
  private val ctx = new ThreadLocal[MemorySegment] {
    override def initialValue(): MemorySegment = {
      val out = Arena.global().allocate(ADDRESS)
      check(ctxCreate.invoke(0, out).asInstanceOf[Int])
      out.get(ADDRESS, 0)
    }
  }

  private def check(status: Int): Unit =
    if (status != 0) throw new RuntimeException(s"gpb200 status $status") // no CPU fallback

  private def requireEuclidean(dist: (Location[Double], Location[Double]) => Double): Unit = {
    // dist is an arbitrary closure in the reference; only Location.euclidean ships (SURVEY 0.2m):
    // probe it on a few pairs and refuse anything else instead of falling back to the CPU.
    val probes = Vector((One(0.0), One(3.0)), (One(-1.5), One(2.25)))
    if (!probes.forall { case (a, b) => dist(a, b) == Location.euclidean(a, b) })
      throw new UnsupportedOperationException("gpb200 supports Location.euclidean only")
  }

  private def soa(xs: Vector[Location[Double]], arena: Arena): (MemorySegment, Int) = {
    val dim = xs.head.toVector.size
    val seg = arena.allocate(JAVA_DOUBLE, xs.size.toLong * dim)
    for ((x, i) <- xs.zipWithIndex; (v, d) <- x.toVector.zipWithIndex)
      seg.setAtIndex(JAVA_DOUBLE, d.toLong * xs.size + i, v)
    (seg, dim)
  }

  private def terms(ps: Vector[KernelParameters], arena: Arena): MemorySegment = {
    val seg = arena.allocate(TERM, ps.size)
    ps.zipWithIndex.foreach { case (p, i) =>
      val (kind, p0, p1) = p match {
        case SquaredExp(h, s) => (0, h, s)
        case White(s)         => (1, s, 0.0)
        case Matern(s, nu, _) => (2, s, nu) // -> GPB_E_UNSUPPORTED
      }
      val o = i * TERM.byteSize()
      seg.set(JAVA_INT, o, kind); seg.set(JAVA_DOUBLE, o + 8, p0); seg.set(JAVA_DOUBLE, o + 16, p1)
    }
    seg
  }

  /** Same signature and currying as GaussianProcess.loglikelihood (GaussianProcess.scala:109-111):
    * the observations are uploaded when the outer function is applied. */
  def loglikelihood(observed: Vector[Data],
                    dist: (Location[Double], Location[Double]) => Double): Parameters => Double = {
    requireEuclidean(dist)
    val arena = Arena.ofShared()
    val (x, dim) = soa(observed.map(_.x), arena)
    val y = arena.allocateFrom(JAVA_DOUBLE, observed.map(_.y): _*)
    val dsOut = arena.allocate(ADDRESS)
    check(dsCreate.invoke(ctx.get, observed.size, dim, x, y, dsOut).asInstanceOf[Int])
    val ds = dsOut.get(ADDRESS, 0)

    { p: Parameters =>
      val a = Arena.ofConfined()
      try {
        val beta = p.meanParameters.toList
        val out = a.allocate(LL)
        check(loglik.invoke(ctx.get, ds, terms(p.kernelParameters, a), p.kernelParameters.size,
          if (beta.isEmpty) MemorySegment.NULL else a.allocateFrom(JAVA_DOUBLE, beta: _*),
          beta.size, 1e-3, out).asInstanceOf[Int])
        if (out.get(JAVA_INT, 28) != 0) // breeze.linalg.cholesky throws on a non-PD matrix
          throw new breeze.linalg.NotConvergedException(breeze.linalg.NotConvergedException.Divergence)
        llValue.invoke(out, 0 /* GPB_FORMULA_AS_PARSED: what :126 computes */).asInstanceOf[Double]
      } finally a.close()
    }
  }

  /** Same signature as Predict.fit (Predict.scala:57-61). */
  def fit(newXs: Vector[Location[Double]], observed: Vector[Data],
          dist: (Location[Double], Location[Double]) => Double,
          p: Parameters): Vector[(Location[Double], Gaussian)] = {
    requireEuclidean(dist)
    val a = Arena.ofConfined()
    try {
      val (x, dim) = soa(observed.map(_.x), a)
      val y = a.allocateFrom(JAVA_DOUBLE, observed.map(_.y): _*)
      val dsOut = a.allocate(ADDRESS)
      check(dsCreate.invoke(ctx.get, observed.size, dim, x, y, dsOut).asInstanceOf[Int])
      val ds = dsOut.get(ADDRESS, 0)
      val (xs, _) = soa(newXs, a)
      val m = newXs.size
      val mean = a.allocate(JAVA_DOUBLE, m); val vari = a.allocate(JAVA_DOUBLE, m)
      val info = a.allocate(JAVA_INT)
      check(predict.invoke(ctx.get, ds, terms(p.kernelParameters, a), p.kernelParameters.size, 1e-6, m,
        xs, mean, vari, info).asInstanceOf[Int])
      dsDestroy.invoke(ds)
      if (info.get(JAVA_INT, 0) != 0)
        throw new breeze.linalg.NotConvergedException(breeze.linalg.NotConvergedException.Divergence)
      newXs.zipWithIndex.map { case (l, i) =>
        (l, Gaussian(mean.getAtIndex(JAVA_DOUBLE, i), math.sqrt(vari.getAtIndex(JAVA_DOUBLE, i))))
      }
    } finally a.close()
  }

  private val samplePlaneH = fn("gpb_sample_plane", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS,
    JAVA_INT, JAVA_DOUBLE, JAVA_INT, ADDRESS, JAVA_DOUBLE, JAVA_DOUBLE, ADDRESS, ADDRESS, ADDRESS))
  private val drawH = fn("gpb_draw", FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT, JAVA_INT, ADDRESS,
    ADDRESS, JAVA_INT, JAVA_DOUBLE, ADDRESS, ADDRESS, ADDRESS, ADDRESS))

  /** Same signature as MeanParameters.samplePlane (MeanParameters.scala:74-79).  The design matrix is
    * built by the reference's own makeDesignMatrix (layout untouched) and handed over column-major. */
  def samplePlane(prior: Gaussian, obs: Vector[Data], xs: Vector[Location[Double]],
                  dist: (Location[Double], Location[Double]) => Double, p: Parameters): MeanParameters = {
    requireEuclidean(dist)
    val a = Arena.ofConfined()
    try {
      val (x, dim) = soa(xs, a)
      val y = a.allocateFrom(JAVA_DOUBLE, obs.map(_.y): _*)
      val dsOut = a.allocate(ADDRESS)
      check(dsCreate.invoke(ctx.get, xs.size, dim, x, y, dsOut).asInstanceOf[Int])
      val ds = dsOut.get(ADDRESS, 0)
      val design = MeanParameters.makeDesignMatrix(xs)            // n x d, column-major == [d][n]
      val d = design.cols
      val z = a.allocateFrom(JAVA_DOUBLE, Seq.fill(d)(Gaussian(0, 1).draw): _*)
      val beta = a.allocate(JAVA_DOUBLE, d); val info = a.allocate(JAVA_INT)
      check(samplePlaneH.invoke(ctx.get, ds, terms(p.kernelParameters, a), p.kernelParameters.size, 1e-6, d,
        a.allocateFrom(JAVA_DOUBLE, design.toArray: _*), prior.mean, prior.variance, z, beta, info).asInstanceOf[Int])
      dsDestroy.invoke(ds)
      if (info.get(JAVA_INT, 0) != 0)
        throw new breeze.linalg.NotConvergedException(breeze.linalg.NotConvergedException.Divergence)
      MeanParameters.plane(breeze.linalg.DenseVector(beta.toArray(JAVA_DOUBLE)))
    } finally a.close()
  }

  /** Same signature as GaussianProcess.draw (GaussianProcess.scala:67-69); Cholesky root. */
  def draw(xs: Vector[Location[Double]], dist: (Location[Double], Location[Double]) => Double,
           p: Parameters): breeze.linalg.DenseVector[Double] = {
    requireEuclidean(dist)
    val a = Arena.ofConfined()
    try {
      val (x, dim) = soa(xs, a)
      val mu = a.allocateFrom(JAVA_DOUBLE, xs.map(MeanFunction.apply(p.meanParameters)): _*)
      val z = a.allocateFrom(JAVA_DOUBLE, Seq.fill(xs.size)(Gaussian(0, 1).draw): _*)
      val out = a.allocate(JAVA_DOUBLE, xs.size); val info = a.allocate(JAVA_INT)
      check(drawH.invoke(ctx.get, xs.size, dim, x, terms(p.kernelParameters, a), p.kernelParameters.size,
        1e-3, mu, z, out, info).asInstanceOf[Int])
      breeze.linalg.DenseVector(out.toArray(JAVA_DOUBLE))
    } finally a.close()
  }
  private val mllGradientH = fn("gpb_mll_gradient", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_INT,
    ADDRESS, JAVA_INT, ADDRESS, JAVA_INT, JAVA_DOUBLE, JAVA_INT, ADDRESS, ADDRESS))

  /** Same signature and currying as GaussianProcess.mllGradient (GaussianProcess.scala:166-169); p holds
    * the unconstrained kernel parameters, as sampleHmc / sampleEhmc pass them.  mode 0 = the formulas
    * the Scala executes, 1 = the analytic gradient in the log parameters. */
  def mllGradient(observed: Vector[Data], dist: (Location[Double], Location[Double]) => Double,
                  mode: Int = 0): Parameters => Array[Double] = {
    requireEuclidean(dist)
    val arena = Arena.ofShared()
    val (x, dim) = soa(observed.map(_.x), arena)
    val y = arena.allocateFrom(JAVA_DOUBLE, observed.map(_.y): _*)
    val dsOut = arena.allocate(ADDRESS)
    check(dsCreate.invoke(ctx.get, observed.size, dim, x, y, dsOut).asInstanceOf[Int])
    val ds = dsOut.get(ADDRESS, 0)

    { p: Parameters =>
      val a = Arena.ofConfined()
      try {
        val beta = p.meanParameters.toList
        val np = p.kernelParameters.map(_.toList.size).sum
        val grad = a.allocate(JAVA_DOUBLE, np); val info = a.allocate(JAVA_INT)
        check(mllGradientH.invoke(ctx.get, ds, 1, terms(p.kernelParameters, a), p.kernelParameters.size,
          if (beta.isEmpty) MemorySegment.NULL else a.allocateFrom(JAVA_DOUBLE, beta: _*),
          beta.size, 1e-3, mode, grad, info).asInstanceOf[Int])
        if (info.get(JAVA_INT, 0) != 0) // kxx not positive definite: the LU of the reference would not notice
          throw new breeze.linalg.NotConvergedException(breeze.linalg.NotConvergedException.Divergence)
        grad.toArray(JAVA_DOUBLE)
      } finally a.close()
    }
  }

  // ---- Mcmc.sample for many chains, device resident (gpb_mh_run_ex), and the multi-GPU gather ------------
  // struct gpb_mh_config (include/gpb200.h): ints and doubles in declaration order, natural alignment
  private val MHCFG = MemoryLayout.structLayout(
    JAVA_INT.withName("struct_size"), JAVA_INT.withName("formula"), JAVA_INT.withName("on_notpd"), JAVA_INT.withName("rng"),
    JAVA_DOUBLE.withName("delta"), JAVA_DOUBLE.withName("nugget"), JAVA_LONG.withName("seed"),
    JAVA_INT.withName("chain0"), JAVA_INT.withName("mean_step"),
    JAVA_DOUBLE.withName("mean_prior_mean"), JAVA_DOUBLE.withName("mean_prior_var"), JAVA_DOUBLE.withName("mean_nugget"),
    JAVA_INT.withName("flush_every"), JAVA_INT.withName("gather"), JAVA_INT.withName("total_chains"), JAVA_INT.withName("small_n"))
  // struct gpb_mh_result { int info, stopped_chains, path, flushes; float device_ms; }
  private val MHRES = MemoryLayout.structLayout(JAVA_INT, JAVA_INT, JAVA_INT, JAVA_INT, JAVA_FLOAT)
  private val mhRunExH = fn("gpb_mh_run_ex", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS, JAVA_INT, JAVA_INT,
    ADDRESS, JAVA_INT, ADDRESS, JAVA_INT, ADDRESS, ADDRESS, ADDRESS, ADDRESS, ADDRESS, ADDRESS, ADDRESS, ADDRESS, ADDRESS))
  private val commIdH = fn("gpb_comm_unique_id", FunctionDescriptor.of(JAVA_INT, ADDRESS))
  private val commInitH = fn("gpb_comm_init_rank", FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT, JAVA_INT, ADDRESS))
  private val commShardH = fn("gpb_comm_shard", FunctionDescriptor.of(JAVA_INT, JAVA_INT, JAVA_INT, JAVA_INT, ADDRESS, ADDRESS))

  case class MhConfig(delta: Double = 0.05, formula: Int = 0, onNotPd: Int = 0, seed: Long = 0L, chain0: Int = 0,
                      meanStep: Boolean = false, priorMean: Gaussian = Gaussian(0, 1), flushEvery: Int = 0,
                      gather: Boolean = false, totalChains: Int = 0)
  case class Chains(states: Array[Double] /* [iters][C][W] */, ll: Array[Double], accept: Array[Int], stop: Array[Int])

  /** The 128-byte communicator id: create it once, hand it to every rank's thread. */
  def commUniqueId(): Array[Byte] = {
    val a = Arena.ofConfined()
    try { val id = a.allocate(128); check(commIdH.invoke(id).asInstanceOf[Int]); id.toArray(JAVA_BYTE) } finally a.close()
  }
  /** One rank = one thread = one ctx = one device (the threading model of SimulatedGp.scala:163-165). */
  def commInitRank(nranks: Int, rank: Int, id: Array[Byte]): Unit = {
    val a = Arena.ofConfined()
    try check(commInitH.invoke(ctx.get, nranks, rank, a.allocateFrom(JAVA_BYTE, id: _*)).asInstanceOf[Int]) finally a.close()
  }
  def commShard(total: Int, nranks: Int, rank: Int): (Int, Int) = {
    val a = Arena.ofConfined()
    try {
      val f = a.allocate(JAVA_INT); val c = a.allocate(JAVA_INT)
      check(commShardH.invoke(total, nranks, rank, f, c).asInstanceOf[Int]); (f.get(JAVA_INT, 0), c.get(JAVA_INT, 0))
    } finally a.close()
  }

  /** Mcmc.sample(ys, priorKern, priorMean, propKern, dist, initP) for `inits.size` chains on this thread's
    * device, with the shipped example's proposal (log-normal random walk, step delta) and inverse-gamma priors
    * `ig` = (shape, scale) per kernel parameter; device Philox streams.  With cfg.gather every rank's call returns
    * the chains of ALL ranks (NCCL all-gather of every ring segment inside the library). */
  def mhRunEx(observed: Vector[Data], dist: (Location[Double], Location[Double]) => Double, inits: Vector[Parameters],
              ig: Vector[(Double, Double)], iters: Int, cfg: MhConfig): Chains = {
    requireEuclidean(dist)
    val a = Arena.ofConfined()
    try {
      val (x, dim) = soa(observed.map(_.x), a)
      val y = a.allocateFrom(JAVA_DOUBLE, observed.map(_.y): _*)
      val dsOut = a.allocate(ADDRESS)
      check(dsCreate.invoke(ctx.get, observed.size, dim, x, y, dsOut).asInstanceOf[Int])
      val ds = dsOut.get(ADDRESS, 0)
      val nterms = inits.head.kernelParameters.size
      val P = inits.head.kernelParameters.map(_.toList.size).sum
      val nbeta = inits.head.meanParameters.toList.size
      val d = if (cfg.meanStep) nbeta else 0
      val c = MHCFG.byteSize(); val m = a.allocate(MHCFG)
      m.set(JAVA_INT, 0, c.toInt); m.set(JAVA_INT, 4, cfg.formula); m.set(JAVA_INT, 8, cfg.onNotPd); m.set(JAVA_INT, 12, 1 /* GPB_RNG_PHILOX */)
      m.set(JAVA_DOUBLE, 16, cfg.delta); m.set(JAVA_DOUBLE, 24, 1e-3); m.set(JAVA_LONG, 32, cfg.seed)
      m.set(JAVA_INT, 40, cfg.chain0); m.set(JAVA_INT, 44, if (cfg.meanStep) 1 else 0)
      m.set(JAVA_DOUBLE, 48, cfg.priorMean.mean); m.set(JAVA_DOUBLE, 56, cfg.priorMean.variance); m.set(JAVA_DOUBLE, 64, 1e-6)
      m.set(JAVA_INT, 72, cfg.flushEvery); m.set(JAVA_INT, 76, if (cfg.gather) 1 else 0); m.set(JAVA_INT, 80, cfg.totalChains); m.set(JAVA_INT, 84, 0)
      val termsAll = a.allocate(TERM, inits.size.toLong * nterms)
      inits.zipWithIndex.foreach { case (p, i) => MemorySegment.copy(terms(p.kernelParameters, a), 0, termsAll, i.toLong * nterms * TERM.byteSize(), nterms * TERM.byteSize()) }
      val beta0 = if (nbeta == 0) MemorySegment.NULL else a.allocateFrom(JAVA_DOUBLE, inits.flatMap(_.meanParameters.toList): _*)
      val design = if (!cfg.meanStep) MemorySegment.NULL
                   else a.allocateFrom(JAVA_DOUBLE, MeanParameters.makeDesignMatrix(observed.map(_.x).distinct).toArray: _*)  // column-major [d][n]
      val C = if (cfg.gather) cfg.totalChains else inits.size
      val W = P + d
      val st = a.allocate(JAVA_DOUBLE, iters.toLong * C * W); val ll = a.allocate(JAVA_DOUBLE, iters.toLong * C)
      val acc = a.allocate(JAVA_INT, iters.toLong * C); val stop = a.allocate(JAVA_INT, C.toLong); val res = a.allocate(MHRES)
      val rc = mhRunExH.invoke(ctx.get, ds, m, inits.size, iters, termsAll, nterms, beta0, nbeta,
        a.allocateFrom(JAVA_DOUBLE, ig.flatMap(t => Seq(t._1, t._2)): _*), MemorySegment.NULL, MemorySegment.NULL, design,
        st, ll, acc, stop, res).asInstanceOf[Int]
      dsDestroy.invoke(ds)
      if (rc == -7) // GPB_E_NOTPD: `cholesky` threw on some chain's thread in the reference
        throw new breeze.linalg.NotConvergedException(breeze.linalg.NotConvergedException.Divergence)
      check(rc)
      Chains(st.toArray(JAVA_DOUBLE), ll.toArray(JAVA_DOUBLE), acc.toArray(JAVA_INT), stop.toArray(JAVA_INT))
    } finally a.close()
  }
}
