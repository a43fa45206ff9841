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
}
