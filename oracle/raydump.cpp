/*
 * oracle/raydump.cpp — TEST INFRASTRUCTURE (golden-vector generator + CPU reference timing).
 *
 * A Mitsuba *utility plugin* (loaded by the reference's own `mtsutil`, src/mitsuba/mtsutil.cpp:300-345)
 * written for this repository and compiled against the UNMODIFIED reference headers/libraries built by
 * oracle/Makefile.ref.  It calls the reference's public API only:
 *
 *   trace   : ShapeKDTree::rayIntersect(ray, its)  (src/librender/skdtree.cpp:112-142)  closest hit
 *             ShapeKDTree::rayIntersect(ray)       (src/librender/skdtree.cpp:207-226)  any hit
 *   bsdf    : BSDF::eval / pdf / sample            (include/mitsuba/render/bsdf.h:215-445)
 *   emitter : Scene::sampleEmitterDirect / pdfEmitterDirect (src/librender/scene.cpp:860-884,981-984)
 *
 * Usage (all files raw little-endian float32/int32, no header):
 *   mtsutil raydump trace   <scene.xml|mesh.ply> <rays.bin> <out.bin> [threads] [repeats]
 *       rays : N x 8  {o.xyz, mint, d.xyz, maxt}
 *       out  : N x 16 {t, u, v, shape(i32), prim(i32), p.xyz, ng.xyz, ns.xyz, any(i32), 0}
 *   mtsutil raydump bsdf    <scene.xml> <in.bin> <out.bin>
 *       in   : N x 12 {shape, wi.xyz, wo.xyz, s1, s2, s3 (the draw from bRec.sampler), 0,0}
 *       out  : N x 16 {eval.rgb, pdf, weight.rgb, wo'.xyz, pdf', eta, sampledType, 0,0,0}
 *   mtsutil raydump emitter <scene.xml> <in.bin> <out.bin>
 *       in   : N x 8  {p.xyz, n.xyz, s1, s2}
 *       out  : N x 12 {value.rgb, pdf, d.xyz, dist, measure, pdfEmitterDirect, n_emitter.z?0, 0}
 * Timing for `trace` is printed as one JSON line on stdout (prefix "RAYDUMP ").
 */
#include <mitsuba/render/util.h>
#include <mitsuba/render/scene.h>
#include <mitsuba/render/skdtree.h>
#include <mitsuba/render/trimesh.h>
#include <mitsuba/render/bsdf.h>
#include <mitsuba/render/emitter.h>
#include <mitsuba/core/plugin.h>
#include <mitsuba/core/fresolver.h>
#include <mitsuba/core/timer.h>
#include <mitsuba/core/filesystem.h>
#include <mitsuba/core/fstream.h>
#include <mitsuba/core/bitmap.h>
#include <mitsuba/render/mipmap.h>
#include "sobolseq.h"            /* src/samplers/sobolseq.h: the Sobol direction numbers (raydump sobol) */
#include <thread>
#include <mitsuba/core/qmc.h>
#include <vector>
#include <chrono>
#include <cstdio>
#include <cstring>

MTS_NAMESPACE_BEGIN

static std::vector<float> readFloats(const char *fn) {
	FILE *f = fopen(fn, "rb");
	if (!f) SLog(EError, "raydump: cannot open \"%s\"", fn);
	fseek(f, 0, SEEK_END);
	size_t sz = (size_t) ftell(f);
	fseek(f, 0, SEEK_SET);
	std::vector<float> v(sz / 4);
	if (fread(v.data(), 4, v.size(), f) != v.size()) SLog(EError, "raydump: short read");
	fclose(f);
	return v;
}

/// Hands the `bsdf` mode's third input sample to plugins that draw from bRec.sampler inside sample() (EUsesSampler:
/// roughdielectric's lobe selection) — so that draw is an input of the golden vector like the other two.
class FixedSampler : public Sampler {
public:
	FixedSampler() : Sampler(Properties()), m_value(0.5f) { }
	void set(Float v) { m_value = v; }
	Float next1D() { return m_value; }
	Point2 next2D() { return Point2(m_value, m_value); }
private:
	Float m_value;
};

static void writeFloats(const char *fn, const std::vector<float> &v) {
	FILE *f = fopen(fn, "wb");
	if (!f) SLog(EError, "raydump: cannot write \"%s\"", fn);
	fwrite(v.data(), 4, v.size(), f);
	fclose(f);
}

static inline float i2f(int32_t i) { float f; memcpy(&f, &i, 4); return f; }

class RayDump : public Utility {
public:
	ref<Scene> m_scene;
	ref<ShapeKDTree> m_kdtree;
	std::vector<const Shape *> m_shapes;

	void load(const char *fn) {
		ref<FileResolver> fileResolver = Thread::getThread()->getFileResolver();
		std::string lower(fn);
		for (auto &c : lower) c = (char) tolower(c);
		bool isPly = lower.size() > 4 && lower.substr(lower.size()-4) == ".ply";
		if (!isPly) {
			fs::path filename = fs::decode_pathstr(fileResolver->resolve(fs::pathstr(fn)));
			ref<FileResolver> frClone = fileResolver->clone();
			frClone->prependPath(fs::encode_pathstr(fs::absolute(filename).parent_path()));
			Thread::getThread()->setFileResolver(frClone);
			m_scene = loadScene(fs::pathstr(fn));
			m_scene->initialize();
			m_kdtree = m_scene->getKDTree();
			/* kd-tree shape order == order of addShape() calls == Scene::m_shapes order */
			const ref_vector<Shape> &shapes = m_scene->getShapes();
			for (size_t i=0; i<shapes.size(); ++i)
				m_shapes.push_back(shapes[i].get());
		} else {
			Properties props("ply");
			props.setString("filename", fn);
			ref<TriMesh> mesh = static_cast<TriMesh *>(PluginManager::getInstance()->
				createObject(MTS_CLASS(TriMesh), props));
			mesh->configure();
			mesh->incRef();
			m_kdtree = new ShapeKDTree();
			m_kdtree->addShape(mesh);
			m_kdtree->build();
			m_shapes.push_back(mesh.get());
		}
	}

	int shapeIndex(const Shape *s) const {
		for (size_t i=0; i<m_shapes.size(); ++i)
			if (m_shapes[i] == s) return (int) i;
		return -1;
	}

	int trace(int argc, char **argv) {
		if (argc < 5) { cout << "trace <scene> <rays> <out> [threads] [repeats]" << endl; return 1; }
		ref<Timer> buildTimer = new Timer();
		load(argv[2]);
		unsigned buildMs = buildTimer->getMilliseconds();
		std::vector<float> rays = readFloats(argv[3]);
		size_t n = rays.size() / 8;
		int nThreads = argc > 5 ? atoi(argv[5]) : 1;
		int repeats = argc > 6 ? atoi(argv[6]) : 1;
		if (nThreads < 1) nThreads = (int) std::thread::hardware_concurrency();
		std::vector<float> out(n * 16, 0.0f);
		const ShapeKDTree *kd = m_kdtree.get();
		Thread *parent = Thread::getThread();
		double bestClosest = 1e30, bestAny = 1e30;
		std::vector<double> allClosest, allAny;

		for (int rep = 0; rep < repeats; ++rep) {
			for (int pass = 0; pass < 2; ++pass) {
				auto t0 = std::chrono::steady_clock::now();
				std::vector<std::thread> threads;
				for (int tid = 0; tid < nThreads; ++tid) {
					threads.emplace_back([&, tid]() {
						Thread *thr = Thread::registerUnmanagedThread("raydump");
						thr->setLogger(parent->getLogger());
						thr->setFileResolver(parent->getFileResolver());
						size_t begin = n * tid / nThreads, end = n * (tid + 1) / nThreads;
						for (size_t i = begin; i < end; ++i) {
							const float *r = &rays[i*8];
							float *o = &out[i*16];
							Ray ray(Point(r[0], r[1], r[2]), Vector(r[4], r[5], r[6]), r[3], r[7], 0.0f);
							if (pass == 0) {
								Intersection its;
								if (kd->rayIntersect(ray, its)) {
									o[0] = its.t; o[1] = its.uv.x; o[2] = its.uv.y;
									o[3] = i2f(shapeIndex(its.shape)); o[4] = i2f((int32_t) its.primIndex);
									o[5] = its.p.x; o[6] = its.p.y; o[7] = its.p.z;
									o[8] = its.geoFrame.n.x; o[9] = its.geoFrame.n.y; o[10] = its.geoFrame.n.z;
									o[11] = its.shFrame.n.x; o[12] = its.shFrame.n.y; o[13] = its.shFrame.n.z;
								} else {
									o[0] = std::numeric_limits<float>::infinity();
									o[3] = i2f(-1); o[4] = i2f(-1);
								}
							} else {
								o[14] = i2f(kd->rayIntersect(ray) ? 1 : 0);
							}
						}
					});
				}
				for (auto &t : threads) t.join();
				double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
				if (pass == 0) { bestClosest = std::min(bestClosest, ms); allClosest.push_back(ms); }
				else { bestAny = std::min(bestAny, ms); allAny.push_back(ms); }
			}
		}
		writeFloats(argv[4], out);
		printf("RAYDUMP {\"rays\": %zu, \"threads\": %d, \"repeats\": %d, \"build_ms\": %u, \"closest_ms\": %.3f, "
			"\"any_ms\": %.3f, \"closest_mrays\": %.4f, \"any_mrays\": %.4f, \"prims\": %u, \"closest_ms_all\": [",
			n, nThreads, repeats, buildMs, bestClosest, bestAny, n / bestClosest / 1e3, n / bestAny / 1e3,
			(unsigned) kd->getPrimitiveCount());
		for (size_t i = 0; i < allClosest.size(); ++i) printf("%s%.3f", i ? ", " : "", allClosest[i]);
		printf("], \"any_ms_all\": [");
		for (size_t i = 0; i < allAny.size(); ++i) printf("%s%.3f", i ? ", " : "", allAny[i]);
		printf("]}\n");
		fflush(stdout);
		return 0;
	}

	/* ShapeKDTree::rayIntersect(ray, t, shape, n, uv) — the "light record" overload (src/librender/skdtree.cpp:144-204):
	   8 floats per ray {t, shape index, n.xyz, uv.xy, 0}; t = +inf and shape = -1 on a miss */
	int light(int argc, char **argv) {
		if (argc < 5) { cout << "light <scene> <rays> <out>" << endl; return 1; }
		load(argv[2]);
		std::vector<float> rays = readFloats(argv[3]);
		size_t n = rays.size() / 8;
		std::vector<float> out(n * 8, 0.0f);
		const ShapeKDTree *kd = m_kdtree.get();
		for (size_t i = 0; i < n; ++i) {
			const float *r = &rays[i*8];
			float *o = &out[i*8];
			Ray ray(Point(r[0], r[1], r[2]), Vector(r[4], r[5], r[6]), r[3], r[7], 0.0f);
			Float t; ConstShapePtr shape = NULL; Normal nrm(0.0f); Point2 uv(0.0f);
			if (kd->rayIntersect(ray, t, shape, nrm, uv)) {
				o[0] = t; o[1] = i2f(shapeIndex(shape));
				o[2] = nrm.x; o[3] = nrm.y; o[4] = nrm.z; o[5] = uv.x; o[6] = uv.y;
			} else {
				o[0] = std::numeric_limits<float>::infinity(); o[1] = i2f(-1);
			}
		}
		writeFloats(argv[4], out);
		return 0;
	}

	/// Scene::rayIntersectAll(ray, its) and Scene::rayIntersectAll(ray) (src/librender/scene.cpp:716-730, 769-792): 4 floats per
	/// ray — t, shape index (kd-tree shapes >= 0, special shape k -> -(2 + k), miss -> -1), any-hit flag, pad
	int all(int argc, char **argv) {
		if (argc < 5) { cout << "all <scene.xml> <rays> <out>" << endl; return 1; }
		load(argv[2]);
		if (!m_scene) { cout << "all: needs a scene file" << endl; return 1; }
		std::vector<float> rays = readFloats(argv[3]);
		size_t n = rays.size() / 8;
		std::vector<float> out(n * 4, 0.0f);
		const ref_vector<Shape> &special = m_scene->getSpecialShapes();
		cout << "special shapes: " << special.size();
		for (size_t k = 0; k < special.size(); ++k) cout << " " << special[k]->getClass()->getName();
		cout << endl;
		for (size_t i = 0; i < n; ++i) {
			const float *r = &rays[i*8];
			float *o = &out[i*4];
			Ray ray(Point(r[0], r[1], r[2]), Vector(r[4], r[5], r[6]), r[3], r[7], 0.0f);
			Intersection its;
			if (m_scene->rayIntersectAll(ray, its)) {
				int idx = shapeIndex(its.shape);
				if (idx < 0)
					for (size_t k = 0; k < special.size(); ++k)
						if (special[k].get() == its.shape) idx = -(2 + (int) k);
				o[0] = its.t; o[1] = i2f(idx);
			} else {
				o[0] = std::numeric_limits<float>::infinity(); o[1] = i2f(-1);
			}
			o[2] = m_scene->rayIntersectAll(ray) ? 1.0f : 0.0f;
		}
		writeFloats(argv[4], out);
		return 0;
	}

	/// The MIP pyramid an `envmap` emitter builds from its image (src/emitters/envmap.cpp:152-178: RGB, half-precision storage,
	/// 2-lobed Lanczos downsampling, repeat / clamp boundaries): floats {levels, then per level: width, height, w*h*3 values}
	int mipmap(int argc, char **argv) {
		if (argc < 4) { cout << "mipmap <image> <out>" << endl; return 1; }
		typedef TSpectrum<half, SPECTRUM_SAMPLES> SpectrumHalf;
		typedef TMIPMap<Spectrum, SpectrumHalf> MIPMap;
		ref<FileStream> fs = new FileStream(fs::pathstr(argv[2]), FileStream::EReadOnly);
		ref<Bitmap> bitmap = new Bitmap(Bitmap::EAuto, fs);
		Properties rfilterProps("lanczos");
		rfilterProps.setInteger("lobes", 2);
		ref<ReconstructionFilter> rfilter = static_cast<ReconstructionFilter *>(PluginManager::getInstance()->createObject(
			MTS_CLASS(ReconstructionFilter), rfilterProps));
		rfilter->configure();
		MIPMap *mip = new MIPMap(bitmap, Bitmap::ERGB, Bitmap::EFloat, rfilter, ReconstructionFilter::ERepeat, ReconstructionFilter::EClamp,
			EEWA, 10.0f, fs::pathstr(), 0, std::numeric_limits<Float>::infinity(), Spectrum::EIlluminant);
		std::vector<float> out;
		out.push_back((float) mip->getLevels());
		for (int l = 0; l < mip->getLevels(); ++l) {
			const MIPMap::Array2DType &a = mip->getArray(l);
			out.push_back((float) a.getSize().x); out.push_back((float) a.getSize().y);
			for (int y = 0; y < a.getSize().y; ++y)
				for (int x = 0; x < a.getSize().x; ++x) {
					Spectrum v(a(x, y));
					Float r, g, b; v.toLinearRGB(r, g, b);
					out.push_back(r); out.push_back(g); out.push_back(b);
				}
		}
		writeFloats(argv[3], out);
		cout << "mipmap: " << mip->getLevels() << " levels from " << bitmap->getWidth() << "x" << bitmap->getHeight() << endl;
		delete mip;
		return 0;
	}

	/// The MIP pyramid a `bitmap` texture builds from an RGB image (src/textures/bitmap.cpp:283-305: float texels, 2-lobed Lanczos
	/// downsampling under the texture's own wrap modes); same output layout as `mipmap`
	int mipmap3(int argc, char **argv) {
		if (argc < 6) { cout << "mipmap3 <image> <out> <wrapU> <wrapV>  (0 repeat, 1 clamp, 2 mirror, 3 zero, 4 one)" << endl; return 1; }
		typedef TMIPMap<Color3, Color3> MIPMap3;
		static const ReconstructionFilter::EBoundaryCondition bc[5] = { ReconstructionFilter::ERepeat, ReconstructionFilter::EClamp,
			ReconstructionFilter::EMirror, ReconstructionFilter::EZero, ReconstructionFilter::EOne };
		ref<FileStream> fs = new FileStream(fs::pathstr(argv[2]), FileStream::EReadOnly);
		ref<Bitmap> bitmap = new Bitmap(Bitmap::EAuto, fs);
		Properties rfilterProps("lanczos");
		rfilterProps.setInteger("lobes", 2);
		ref<ReconstructionFilter> rfilter = static_cast<ReconstructionFilter *>(PluginManager::getInstance()->createObject(
			MTS_CLASS(ReconstructionFilter), rfilterProps));
		rfilter->configure();
		ref<MIPMap3> mip = new MIPMap3(bitmap, Bitmap::ERGB, Bitmap::EFloat, rfilter, bc[atoi(argv[4])], bc[atoi(argv[5])], EEWA, 20.0f, fs::pathstr(), 0);
		std::vector<float> out;
		out.push_back((float) mip->getLevels());
		for (int l = 0; l < mip->getLevels(); ++l) {
			const MIPMap3::Array2DType &a = mip->getArray(l);
			out.push_back((float) a.getSize().x); out.push_back((float) a.getSize().y);
			for (int y = 0; y < a.getSize().y; ++y)
				for (int x = 0; x < a.getSize().x; ++x) {
					const Color3 v = a(x, y);
					out.push_back(v[0]); out.push_back(v[1]); out.push_back(v[2]);
				}
		}
		writeFloats(argv[3], out);
		cout << "mipmap3: " << mip->getLevels() << " levels from " << bitmap->getWidth() << "x" << bitmap->getHeight() << endl;
		return 0;
	}

	/// The tables of the `sobol` sampler (src/samplers/sobolseq.h:30-36) as raw uint32 words: 1024 x 52 matrices32, then the 52-column
	/// rows of vdc_sobol_matrices and vdc_sobol_matrices_inv for m = 1..16 as (lo, hi) pairs
	int sobolTables(int argc, char **argv) {
		if (argc < 3) { cout << "sobol <out>" << endl; return 1; }
		std::vector<uint32_t> out;
		for (uint32_t i = 0; i < sobol::Matrices::num_dimensions * sobol::Matrices::size; ++i) out.push_back(sobol::Matrices::matrices32[i]);
		for (int m = 0; m < 16; ++m) for (int c = 0; c < 52; ++c) { uint64_t v = sobol::Matrices::vdc_sobol_matrices[m][c]; out.push_back((uint32_t) v); out.push_back((uint32_t) (v >> 32)); }
		for (int m = 0; m < 16; ++m) for (int c = 0; c < 52; ++c) { uint64_t v = sobol::Matrices::vdc_sobol_matrices_inv[m][c]; out.push_back((uint32_t) v); out.push_back((uint32_t) (v >> 32)); }
		FILE *f = fopen(argv[2], "wb");
		if (!f) return 1;
		fwrite(out.data(), 4, out.size(), f); fclose(f);
		cout << "sobol: " << out.size() << " words" << endl;
		return 0;
	}

	/// The triangle meshes of a scene as the reference's loaders (ply.cpp, TriMesh::configure / computeNormals) leave them:
	/// per mesh shape the words {V, T, hasNormals, hasUV}, 3V position floats, 3V normal floats (if any), 2V uv floats (if any),
	/// 3T index words.  Prints the wall time of loadScene + initialize (the host parse the device decoder replaces).
	int mesh(int argc, char **argv) {
		if (argc < 4) { cout << "mesh <scene> <out>" << endl; return 1; }
		auto t0 = std::chrono::steady_clock::now();
		load(argv[2]);
		double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
		FILE *f = fopen(argv[3], "wb");
		if (!f) return 1;
		size_t meshes = 0;
		for (const Shape *s : m_shapes) {
			if (!s->getClass()->derivesFrom(MTS_CLASS(TriMesh))) continue;
			const TriMesh *m = static_cast<const TriMesh *>(s);
			uint32_t hdr[4] = { (uint32_t) m->getVertexCount(), (uint32_t) m->getTriangleCount(), m->hasVertexNormals() ? 1u : 0u, m->hasVertexTexcoords() ? 1u : 0u };
			fwrite(hdr, 4, 4, f);
			fwrite(m->getVertexPositions(), sizeof(Point), hdr[0], f);
			if (hdr[2]) fwrite(m->getVertexNormals(), sizeof(Normal), hdr[0], f);
			if (hdr[3]) fwrite(m->getVertexTexcoords(), sizeof(Point2), hdr[0], f);
			fwrite(m->getTriangles(), sizeof(Triangle), hdr[1], f);
			++meshes;
		}
		fclose(f);
		cout << "mesh: " << meshes << " meshes, load " << secs << " s" << endl;
		return 0;
	}

	/// Wall time of the reference's PLY loader alone (PLYLoader constructor + TriMesh::configure), no kd-tree: the host parse
	/// the device decoder replaces, timed for bench reports
	int plyload(int argc, char **argv) {
		if (argc < 3) { cout << "plyload <file.ply>..." << endl; return 1; }
		double total = 0; size_t tris = 0;
		for (int i = 2; i < argc; ++i) {
			auto t0 = std::chrono::steady_clock::now();
			std::string fn(argv[i]);
			Properties props(fn.size() > 11 && fn.substr(fn.size() - 11) == ".serialized" ? "serialized" : "ply");
			props.setString("filename", argv[i]);
			ref<TriMesh> mesh = static_cast<TriMesh *>(PluginManager::getInstance()->createObject(MTS_CLASS(TriMesh), props));
			mesh->configure();
			total += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
			tris += mesh->getTriangleCount();
		}
		cout << "plyload: " << tris << " triangles in " << total << " s" << endl;
		return 0;
	}

	/// radicalInverse2Single / sobol2Single (include/mitsuba/core/qmc.h) on (n, scramble0, scramble1) triples of uint32 words
	int qmc(int argc, char **argv) {
		if (argc < 4) { cout << "qmc <in> <out>" << endl; return 1; }
		std::vector<float> in = readFloats(argv[2]);
		std::vector<float> out;
		for (size_t i = 0; i + 2 < in.size(); i += 3) {
			uint32_t w[3]; memcpy(w, &in[i], 12);
			out.push_back(radicalInverse2Single(w[0], w[1]));
			out.push_back(sobol2Single(w[0], w[2]));
		}
		writeFloats(argv[3], out);
		return 0;
	}

	/// Scene::evalTransmittance on segments {p1[3], flags (bit 0 p1OnSurface, bit 1 p2OnSurface), p2[3], medium index or -1} (8 floats
	/// each, the two integers as float values); media in the order of Scene::getMedia().  Output: 3 floats per segment.
	int transmittance(int argc, char **argv) {
		if (argc < 5) { cout << "transmittance <scene> <in> <out> [maxInteractions]" << endl; return 1; }
		load(argv[2]);
		std::vector<float> in = readFloats(argv[3]);
		const int maxInteractions = argc > 5 ? atoi(argv[5]) : -1;
		const ref_vector<Medium> &media = m_scene->getMedia();
		std::vector<float> out;
		out.reserve(in.size() / 8 * 3);
		auto t0 = std::chrono::steady_clock::now();
		for (size_t i = 0; i + 7 < in.size(); i += 8) {
			const int flags = (int) in[i+3], mi = (int) in[i+7];
			const Medium *medium = mi >= 0 && mi < (int) media.size() ? media[mi].get() : NULL;
			int interactions = maxInteractions;
			Spectrum T = m_scene->evalTransmittance(Point(in[i], in[i+1], in[i+2]), (flags & 1) != 0, Point(in[i+4], in[i+5], in[i+6]), (flags & 2) != 0,
				0.0f, medium, interactions, NULL);
			Float r, g, b; T.toLinearRGB(r, g, b);
			out.push_back(r); out.push_back(g); out.push_back(b);
		}
		const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
		writeFloats(argv[4], out);
		cout << "transmittance: " << out.size() / 3 << " segments, " << media.size() << " media" << endl;
		cout << "transmittance_seconds: " << secs << " (1 thread)" << endl;
		return 0;
	}

	int bsdf(int argc, char **argv) {
		if (argc < 5) { cout << "bsdf <scene> <in> <out>" << endl; return 1; }
		load(argv[2]);
		std::vector<float> in = readFloats(argv[3]);
		size_t n = in.size() / 12;
		std::vector<float> out(n * 16, 0.0f);
		ref<FixedSampler> fixedSampler = new FixedSampler();
		for (size_t i = 0; i < n; ++i) {
			const float *r = &in[i*12];
			float *o = &out[i*16];
			const Shape *shape = m_shapes.at((size_t) r[0]);
			const BSDF *bsdf = shape->getBSDF();
			Intersection its;
			its.p = Point(0.0f); its.t = 1.0f; its.shape = shape;
			its.geoFrame = Frame(Normal(0, 0, 1)); its.shFrame = its.geoFrame;
			its.uv = Point2(0.5f); its.hasUVPartials = false; its.time = 0; its.instance = NULL;
			its.dpdu = Vector(1, 0, 0); its.dpdv = Vector(0, 1, 0);
			Vector wi(r[1], r[2], r[3]), wo(r[4], r[5], r[6]);
			its.wi = wi;
			{
				BSDFSamplingRecord bRec(its, wi, wo, ERadiance);
				Spectrum f = bsdf->eval(bRec);
				Float pdf = bsdf->pdf(bRec);
				Float rr, gg, bb; f.toLinearRGB(rr, gg, bb);
				o[0] = rr; o[1] = gg; o[2] = bb; o[3] = pdf;
			}
			{
				fixedSampler->set(r[9]);
				BSDFSamplingRecord bRec(its, fixedSampler.get(), ERadiance);
				bRec.wi = wi;
				Float pdf = 0;
				Spectrum w = bsdf->sample(bRec, pdf, Point2(r[7], r[8]));
				Float rr, gg, bb; w.toLinearRGB(rr, gg, bb);
				o[4] = rr; o[5] = gg; o[6] = bb;
				o[7] = bRec.wo.x; o[8] = bRec.wo.y; o[9] = bRec.wo.z;
				o[10] = pdf; o[11] = bRec.eta; o[12] = (float) bRec.sampledType;
			}
		}
		writeFloats(argv[4], out);
		return 0;
	}

	int emitter(int argc, char **argv) {
		if (argc < 5) { cout << "emitter <scene> <in> <out>" << endl; return 1; }
		load(argv[2]);
		std::vector<float> in = readFloats(argv[3]);
		size_t n = in.size() / 8;
		std::vector<float> out(n * 12, 0.0f);
		for (size_t i = 0; i < n; ++i) {
			const float *r = &in[i*8];
			float *o = &out[i*12];
			DirectSamplingRecord dRec(Point(r[0], r[1], r[2]), 0.0f);
			dRec.refN = Normal(r[3], r[4], r[5]);
			Spectrum value = m_scene->sampleEmitterDirect(dRec, Point2(r[6], r[7]), false);
			Float rr, gg, bb; value.toLinearRGB(rr, gg, bb);
			o[0] = rr; o[1] = gg; o[2] = bb; o[3] = dRec.pdf;
			o[4] = dRec.d.x; o[5] = dRec.d.y; o[6] = dRec.d.z; o[7] = dRec.dist;
			o[8] = (float) dRec.measure;
			o[9] = (dRec.pdf > 0 && !value.isZero()) ? m_scene->pdfEmitterDirect(dRec) : 0.0f;
		}
		writeFloats(argv[4], out);
		return 0;
	}

	int run(int argc, char **argv) {
		if (argc < 2) { cout << "raydump trace|light|all|mipmap|bsdf|emitter ..." << endl; return 1; }
		if (!strcmp(argv[1], "trace")) return trace(argc, argv);
		if (!strcmp(argv[1], "light")) return light(argc, argv);
		if (!strcmp(argv[1], "all")) return all(argc, argv);
		if (!strcmp(argv[1], "mipmap")) return mipmap(argc, argv);
		if (!strcmp(argv[1], "mipmap3")) return mipmap3(argc, argv);
		if (!strcmp(argv[1], "sobol")) return sobolTables(argc, argv);
		if (!strcmp(argv[1], "mesh")) return mesh(argc, argv);
		if (!strcmp(argv[1], "qmc")) return qmc(argc, argv);
		if (!strcmp(argv[1], "transmittance")) return transmittance(argc, argv);
		if (!strcmp(argv[1], "plyload")) return plyload(argc, argv);
		if (!strcmp(argv[1], "bsdf")) return bsdf(argc, argv);
		if (!strcmp(argv[1], "emitter")) return emitter(argc, argv);
		cout << "unknown mode" << endl;
		return 1;
	}

	MTS_DECLARE_UTILITY()
};

MTS_EXPORT_UTILITY(RayDump, "Dump hit records / BSDF / emitter values through the reference API")
MTS_NAMESPACE_END
