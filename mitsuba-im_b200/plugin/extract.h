// extract.h — loaded Mitsuba Scene -> flat arrays of the C ABI (include/mtsb200.h).
//
// Host side of the drop-in: reads the reference's own object graph through its public accessors
// (include/mitsuba/render/scene.h:1065-1100, trimesh.h, ConfigurableObject::getProperties) and re-derives plugin
// parameters the way the plugin constructors do (file:line cited per item).  Anything outside the GPU subset is
// rejected with Log(EError) — there is no CPU fallback.
#pragma once
#include <mitsuba/render/scene.h>
#include <mitsuba/render/trimesh.h>
#include <mitsuba/render/bsdf.h>
#include <mitsuba/render/emitter.h>
#include <mitsuba/render/sensor.h>
#include <mitsuba/render/film.h>
#include <mitsuba/render/sampler.h>
#include <mitsuba/core/fresolver.h>
#include <mitsuba/core/plugin.h>
#include <mitsuba/core/fstream.h>
#include <mitsuba/core/mmap.h>
#include <mitsuba/core/bitmap.h>
#include <mitsuba/render/mipmap.h>
#include "microfacet.h"          /* src/bsdfs/microfacet.h: MicrofacetDistribution(props) reads alpha / distribution / sampleVisible */
#include "rtrans.h"              /* src/bsdfs/rtrans.h: RoughTransmittance, header-only, as roughplastic.cpp includes it */
#include "sobolseq.h"            /* src/samplers/sobolseq.h: sobol::Matrices (the tables are linked in from sobolseq.cpp) */
#include "ior.h"                 /* src/bsdfs/ior.h: lookupIOR(), header-only, as the BSDF plugins include it */
#include "../../include/mtsb200.h"
#include <map>

MTS_NAMESPACE_BEGIN

#define MTSB_CALL(expr) do { int rc__ = (expr); if (rc__ < 0) SLog(EError, "libmtsb200: %s failed (%d): %s", #expr, rc__, mtsb_last_error()); } while (0)

static inline void toRGB(const Spectrum &s, float out[3]) {
	Float r, g, b; s.toLinearRGB(r, g, b);
	out[0] = (float) r; out[1] = (float) g; out[2] = (float) b;
}

static inline void matrixToArray(const Matrix4x4 &m, float out[16]) {
	for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) out[r*4+c] = (float) m(r, c);
}

static inline Intersection dummyIntersection() {
	Intersection its;
	its.p = Point(0.0f); its.t = 1.0f; its.shape = NULL;
	its.geoFrame = Frame(Normal(0, 0, 1)); its.shFrame = its.geoFrame;
	its.uv = Point2(0.5f); its.hasUVPartials = false; its.time = 0; its.instance = NULL;
	its.dpdu = Vector(1, 0, 0); its.dpdv = Vector(0, 1, 0); its.wi = Vector(0, 0, 1);
	return its;
}

/// BSDF plugin -> mtsb_bsdf_desc.  Class names are the reference's RTTI names.
/// The reference's own RoughTransmittance, opened up far enough to read the reduced table back
struct RoughTransmittanceAccess : public RoughTransmittance {
	RoughTransmittanceAccess(MicrofacetDistribution::EType type) : RoughTransmittance(type) { }
	const Float *trans() const { return m_trans; }
	size_t thetaSamples() const { return m_thetaSamples; }
};

/// Switches a render to the `sobol` sampler when the scene uses it (src/samplers/sobol.cpp); other samplers keep the independent
/// stream (or the shared counter stream of oracle/ctrsampler.cpp).  Low-discrepancy samplers the device does not have are refused.
static void configureDeviceSampler(mtsb_render *r, const Sampler *sampler) {
	const std::string cls = sampler->getClass()->getName();
	if (cls == "SobolSampler") {
		if (sampler->getProperties().getSize("scramble", 0) != 0) SLog(EError, "gpupath: the sobol sampler with a scramble value is outside the GPU subset");
		std::vector<uint32_t> t;
		t.reserve(MTSB_SOBOL_TABLE_WORDS);
		for (uint32_t i = 0; i < sobol::Matrices::num_dimensions * sobol::Matrices::size; ++i) t.push_back(sobol::Matrices::matrices32[i]);
		for (int m = 0; m < 16; ++m) for (int c = 0; c < 52; ++c) { const uint64_t v = sobol::Matrices::vdc_sobol_matrices[m][c]; t.push_back((uint32_t) v); t.push_back((uint32_t) (v >> 32)); }
		for (int m = 0; m < 16; ++m) for (int c = 0; c < 52; ++c) { const uint64_t v = sobol::Matrices::vdc_sobol_matrices_inv[m][c]; t.push_back((uint32_t) v); t.push_back((uint32_t) (v >> 32)); }
		MTSB_CALL(mtsb_render_set_sobol(r, t.data(), t.size()));
	} else if (cls == "CounterSampler" && sampler->getProperties().getInteger("ldDimension", 0) > 0) {
		// the oracle's counter sampler in its low-discrepancy mode (oracle/ctrsampler.cpp): the device's ldsampler numbers
		MTSB_CALL(mtsb_render_set_ldsampler(r, sampler->getProperties().getInteger("ldDimension", 0), sampler->getSampleCount()));
	} else if (cls == "LowDiscrepancySampler") {
		// ldsampler.cpp:78-87: `dimension` (default 4); getSampleCount() is already rounded up to a power of two
		const int dimension = sampler->getProperties().getInteger("dimension", 4);
		if (dimension > 15) SLog(EError, "gpupath: ldsampler with dimension %d (the device keeps 15)", dimension);
		MTSB_CALL(mtsb_render_set_ldsampler(r, dimension, sampler->getSampleCount()));
	} else if (cls == "HaltonSampler" || cls == "HammersleySampler" || cls == "StratifiedSampler") {
		SLog(EError, "gpupath: sampler \"%s\" is outside the GPU subset (independent, sobol, ldsampler)", cls.c_str());
	}
}

/// A `bitmap` texture as the device takes it: the MIP pyramid the texture's own constructor builds (src/textures/bitmap.cpp:248-305),
/// rebuilt here with the reference's Bitmap / MIPMap classes from the texture's file, plus the Texture2D parameters.
struct TextureData {
	bool present;
	std::vector<int32_t> sizes; std::vector<std::vector<float> > levels;
	int wrapU, wrapV, filterType; float maxAnisotropy, uvScale[2], uvOffset[2];
	TextureData() : present(false) { }
};

static int parseWrapMode(const std::string &m) {
	if (m == "repeat") return 0; if (m == "clamp") return 1; if (m == "mirror") return 2;
	if (m == "zero" || m == "black") return 3; if (m == "one" || m == "white") return 4;
	SLog(EError, "gpupath: unknown texture wrap mode \"%s\"", m.c_str());
	return 0;
}

static void describeBitmapTexture(const Texture *tex, TextureData &td) {
	typedef TMIPMap<Color3, Color3> MIPMap3;
	static const ReconstructionFilter::EBoundaryCondition bc[5] = { ReconstructionFilter::ERepeat, ReconstructionFilter::EClamp,
		ReconstructionFilter::EMirror, ReconstructionFilter::EZero, ReconstructionFilter::EOne };
	const Properties &props = tex->getProperties();
	if (!props.hasProperty("filename") || !props.getString("channel", "").empty())
		SLog(EError, "gpupath: bitmap textures from raw data or a single channel are outside the GPU subset");
	if (props.getString("coordinates", "uv") != "uv") SLog(EError, "gpupath: only UV texture coordinates are supported");
	const std::string wrap = props.getString("wrapMode", "repeat"), filter = to_lower_copy(props.getString("filterType", "ewa"));
	td.wrapU = parseWrapMode(props.getString("wrapModeU", wrap)); td.wrapV = parseWrapMode(props.getString("wrapModeV", wrap));
	td.filterType = filter == "ewa" ? 3 : (filter == "trilinear" ? 2 : (filter == "bilinear" ? 1 : (filter == "nearest" ? 0 : -1)));
	if (td.filterType < 0) SLog(EError, "gpupath: unknown texture filter type \"%s\"", filter.c_str());
	td.maxAnisotropy = td.filterType == 3 ? (float) props.getFloat("maxAnisotropy", 20) : 1.0f;
	const Float uvscale = props.getFloat("uvscale", 1.0f);
	td.uvScale[0] = (float) props.getFloat("uscale", uvscale); td.uvScale[1] = (float) props.getFloat("vscale", uvscale);
	td.uvOffset[0] = (float) props.getFloat("uoffset", 0.0f); td.uvOffset[1] = (float) props.getFloat("voffset", 0.0f);
	fs::pathstr filename = Thread::getThread()->getFileResolver()->resolve(fs::pathstr(props.getString("filename")));
	ref<FileStream> stream = new FileStream(filename, FileStream::EReadOnly);
	ref<Bitmap> bitmap = new Bitmap(Bitmap::EAuto, stream);
	const Float gamma = props.getFloat("gamma", 0);
	if (gamma != 0) bitmap->setGamma(gamma);
	if (bitmap->getPixelFormat() != Bitmap::ERGB && bitmap->getPixelFormat() != Bitmap::ERGBA)
		SLog(EError, "gpupath: only RGB bitmap textures are inside the GPU subset");
	Properties rfilterProps("lanczos");
	rfilterProps.setInteger("lobes", 2);
	ref<ReconstructionFilter> rfilter = static_cast<ReconstructionFilter *>(PluginManager::getInstance()->createObject(
		MTS_CLASS(ReconstructionFilter), rfilterProps));
	rfilter->configure();
	ref<MIPMap3> mip = new MIPMap3(bitmap, Bitmap::ERGB, Bitmap::EFloat, rfilter, bc[td.wrapU], bc[td.wrapV],
		(EMIPFilterType) td.filterType, td.maxAnisotropy, fs::pathstr(), 0);
	const int levels = mip->getLevels();
	td.sizes.resize(2 * levels); td.levels.resize(levels);
	for (int l = 0; l < levels; ++l) {
		const MIPMap3::Array2DType &a = mip->getArray(l);
		td.sizes[2 * l] = a.getSize().x; td.sizes[2 * l + 1] = a.getSize().y;
		td.levels[l].resize((size_t) a.getSize().x * a.getSize().y * 3);
		float *dst = td.levels[l].data();
		for (int y = 0; y < a.getSize().y; ++y)
			for (int x = 0; x < a.getSize().x; ++x, dst += 3) { const Color3 v = a(x, y); dst[0] = v[0]; dst[1] = v[1]; dst[2] = v[2]; }
	}
	td.present = true;
}

static bool describeBSDF(const BSDF *bsdf, mtsb_bsdf_desc &d, std::vector<float> *table = NULL, TextureData *texture = NULL) {
	memset(&d, 0, sizeof(d));
	const std::string cls = bsdf->getClass()->getName();
	const Properties &props = bsdf->getProperties();
	if (bsdf->usesRayDifferentials() && cls != "SmoothDiffuse" && cls != "SmoothPlastic" && cls != "TwoSidedBRDF")
		SLog(EError, "gpupath: BSDF \"%s\" uses textures that need ray differentials — outside the GPU subset", cls.c_str());
	Intersection its = dummyIntersection();
	if (cls == "SmoothDiffuse") {                                   // src/bsdfs/diffuse.cpp:80-100
		d.type = MTSB_BSDF_DIFFUSE;
		toRGB(bsdf->getDiffuseReflectance(its), d.p);
		if (bsdf->usesRayDifferentials()) {
			// a textured reflectance.  SmoothDiffuse keeps it in `private: ref<Texture> m_reflectance` (diffuse.cpp:188), its only data
			// member, behind no accessor: read through a layout mirror, checked by RTTI (as for twosided below)
			struct Mirror { char base[sizeof(BSDF)]; const Texture *reflectance; };
			const Texture *tex = reinterpret_cast<const Mirror *>(static_cast<const BSDF *>(bsdf))->reflectance;
			if (!tex || !tex->getClass()->derivesFrom(MTS_CLASS(Texture)) || tex->getClass()->getName() != "BitmapTexture")
				SLog(EError, "gpupath: the reflectance texture of this diffuse BSDF is outside the GPU subset (bitmap; a texture brighter than 1 gets wrapped in a scaling texture): %s",
					bsdf->toString().substr(0, 300).c_str());
			if (!texture) SLog(EError, "gpupath: a textured BSDF inside an adapter is outside the GPU subset");
			describeBitmapTexture(tex, *texture);
		}
	} else if (cls == "SmoothDielectric") {                         // src/bsdfs/dielectric.cpp:150-165
		d.type = MTSB_BSDF_DIELECTRIC;
		d.p[0] = (float) bsdf->getEta();
		toRGB(props.getSpectrum("specularReflectance", Spectrum(1.0f)), d.p + 1);
		toRGB(props.getSpectrum("specularTransmittance", Spectrum(1.0f)), d.p + 4);
	} else if (cls == "SmoothPlastic") {                            // src/bsdfs/plastic.cpp:148-215
		d.type = MTSB_BSDF_PLASTIC;
		const Float intIOR = lookupIOR(props, "intIOR", "polypropylene"), extIOR = lookupIOR(props, "extIOR", "air");
		const Float eta = intIOR / extIOR;
		Spectrum diff = props.getSpectrum("diffuseReflectance", Spectrum(0.5f));
		const Spectrum spec = props.getSpectrum("specularReflectance", Spectrum(1.0f));
		const Texture *texturedDiffuse = NULL;
		{
			// textures arrive through addChild, not through the properties: SmoothPlastic keeps them in private members behind no
			// accessor (plastic.cpp:475-480) — read through a layout mirror, checked against the eta computed above and by RTTI
			struct Mirror { char base[sizeof(BSDF)]; Float fdrInt, fdrExt, eta, invEta2; const Texture *diffuse; const Texture *specular; Float ssw; bool nonlinear; };
			const Mirror *m = reinterpret_cast<const Mirror *>(static_cast<const BSDF *>(bsdf));
			if (m->eta != eta || !m->diffuse || !m->specular || !m->diffuse->getClass()->derivesFrom(MTS_CLASS(Texture)) || !m->specular->getClass()->derivesFrom(MTS_CLASS(Texture)))
				SLog(EError, "gpupath: SmoothPlastic does not have the layout of src/bsdfs/plastic.cpp:475-480");
			if (!m->specular->isConstant()) SLog(EError, "gpupath: a textured specularReflectance is outside the GPU subset");
			if (!m->diffuse->isConstant()) {
				if (m->diffuse->getClass()->getName() != "BitmapTexture")
					SLog(EError, "gpupath: the diffuseReflectance texture of this plastic BSDF is outside the GPU subset (bitmap; a texture brighter than 1 gets wrapped in a scaling texture): %s",
						bsdf->toString().substr(0, 300).c_str());
				if (!texture) SLog(EError, "gpupath: a textured BSDF inside an adapter is outside the GPU subset");
				describeBitmapTexture(m->diffuse, *texture);
				diff = m->diffuse->getAverage();                       // plastic.cpp:202-205 weighs the lobes by the texture's average
				texturedDiffuse = m->diffuse;
			}
		}
		d.p[0] = (float) eta;
		toRGB(diff, d.p + 1); toRGB(spec, d.p + 4);
		// the record's own colour is replaced at every hit; the value at the dummy intersection keeps selfCheckBSDF meaningful
		if (texturedDiffuse) toRGB(texturedDiffuse->eval(its), d.p + 1);
		d.p[7] = (float) fresnelDiffuseReflectance(1 / eta, false);
		d.p[8] = (float) fresnelDiffuseReflectance(eta, false);
		const Float dAvg = diff.getLuminance(), sAvg = spec.getLuminance();
		d.p[9] = (float) (sAvg / (dAvg + sAvg));
		d.p[10] = props.getBoolean("nonlinear", false) ? 1.0f : 0.0f;
	} else if (cls == "RoughConductor") {                           // src/bsdfs/roughconductor.cpp:171-203
		d.type = MTSB_BSDF_ROUGHCONDUCTOR;
		ref<FileResolver> fResolver = Thread::getThread()->getFileResolver();
		const std::string materialName = props.getString("material", "Cu");
		Spectrum intEta, intK;
		if (to_lower_copy(materialName) == "none") { intEta = Spectrum(0.0f); intK = Spectrum(1.0f); }
		else {
			intEta.fromContinuousSpectrum(InterpolatedSpectrum(fResolver->resolve(fs::pathstr("data/ior/" + materialName + ".eta.spd"))));
			intK.fromContinuousSpectrum(InterpolatedSpectrum(fResolver->resolve(fs::pathstr("data/ior/" + materialName + ".k.spd"))));
		}
		const Float extEta = lookupIOR(props, "extEta", "air");
		Float alphaU = 0.1f, alphaV = 0.1f;                         // MicrofacetDistribution(props), src/bsdfs/microfacet.h:84-128
		if (props.hasProperty("alpha")) alphaU = alphaV = props.getFloat("alpha");
		else if (props.hasProperty("alphaU") || props.hasProperty("alphaV")) { alphaU = props.getFloat("alphaU"); alphaV = props.getFloat("alphaV"); }
		const std::string distr = to_lower_copy(props.getString("distribution", "beckmann"));
		if (distr != "beckmann" && distr != "ggx")
			SLog(EError, "gpupath: roughconductor distribution \"%s\" is outside the GPU subset (beckmann, ggx)", distr.c_str());
		d.p[0] = (float) alphaU; d.p[1] = (float) alphaV;
		toRGB(props.getSpectrum("eta", intEta) / extEta, d.p + 2);
		toRGB(props.getSpectrum("k", intK) / extEta, d.p + 5);
		toRGB(props.getSpectrum("specularReflectance", Spectrum(1.0f)), d.p + 8);
		d.p[11] = distr == "ggx" ? 1.0f : 0.0f;
		d.p[12] = props.getBoolean("sampleVisible", true) ? 1.0f : 0.0f;
	} else if (cls == "RoughDielectric") {                          // src/bsdfs/roughdielectric.cpp:186-214
		d.type = MTSB_BSDF_ROUGHDIELECTRIC;
		Float alphaU = 0.1f, alphaV = 0.1f;                         // MicrofacetDistribution(props), src/bsdfs/microfacet.h:84-128
		if (props.hasProperty("alpha")) alphaU = alphaV = props.getFloat("alpha");
		else if (props.hasProperty("alphaU") || props.hasProperty("alphaV")) { alphaU = props.getFloat("alphaU"); alphaV = props.getFloat("alphaV"); }
		const std::string distr = to_lower_copy(props.getString("distribution", "beckmann"));
		if (distr != "beckmann" && distr != "ggx")
			SLog(EError, "gpupath: roughdielectric distribution \"%s\" is outside the GPU subset (beckmann, ggx)", distr.c_str());
		d.p[0] = (float) alphaU; d.p[1] = (float) alphaV;
		d.p[2] = (float) bsdf->getEta();
		toRGB(props.getSpectrum("specularReflectance", Spectrum(1.0f)), d.p + 3);
		toRGB(props.getSpectrum("specularTransmittance", Spectrum(1.0f)), d.p + 6);
		d.p[11] = distr == "ggx" ? 1.0f : 0.0f;
		d.p[12] = props.getBoolean("sampleVisible", true) ? 1.0f : 0.0f;
	} else if (cls == "RoughPlastic") {                             // src/bsdfs/roughplastic.cpp:199-303
		d.type = MTSB_BSDF_ROUGHPLASTIC;
		const Float intIOR = lookupIOR(props, "intIOR", "polypropylene"), extIOR = lookupIOR(props, "extIOR", "air");
		const Float eta = intIOR / extIOR;
		MicrofacetDistribution distr(props);
		if (distr.isAnisotropic()) SLog(EError, "gpupath: anisotropic roughplastic is not supported by the plugin either");
		if (props.hasProperty("alpha") && props.getType("alpha") != Properties::EFloat)
			SLog(EError, "gpupath: roughplastic with a textured alpha is outside the GPU subset");
		const Spectrum diff = props.getSpectrum("diffuseReflectance", Spectrum(0.5f));
		const Spectrum spec = props.getSpectrum("specularReflectance", Spectrum(1.0f));
		d.p[0] = (float) distr.getAlpha(); d.p[1] = (float) eta;
		toRGB(diff, d.p + 2); toRGB(spec, d.p + 5);
		const Float dAvg = diff.getLuminance(), sAvg = spec.getLuminance();
		d.p[8] = (float) (sAvg / (dAvg + sAvg));
		// the reductions RoughPlastic::configure performs (:283-303), by the reference's own class
		ref<RoughTransmittanceAccess> ext = new RoughTransmittanceAccess(distr.getType());
		ext->checkEta(eta); ext->checkAlpha(distr.getAlpha());
		ref<RoughTransmittance> internal = ext->clone();
		ext->setEta(eta); internal->setEta(1 / eta);
		ext->setAlpha(distr.getAlpha());
		d.p[9] = (float) (1 - internal->evalDiffuse(distr.getAlpha()));
		d.p[10] = props.getBoolean("nonlinear", false) ? 1.0f : 0.0f;
		d.p[11] = distr.getType() == MicrofacetDistribution::EGGX ? 1.0f : 0.0f;
		if (distr.getType() != MicrofacetDistribution::EGGX && distr.getType() != MicrofacetDistribution::EBeckmann)
			SLog(EError, "gpupath: roughplastic distribution outside the GPU subset (beckmann, ggx)");
		d.p[12] = distr.getSampleVisible() ? 1.0f : 0.0f;
		if (table) table->assign(ext->trans(), ext->trans() + ext->thetaSamples());
	} else if (cls == "Null") {                                     // src/bsdfs/null.cpp:33-38: index-matched medium boundary
		d.type = MTSB_BSDF_PASSTHROUGH;
	} else if (cls == "SmoothConductor") {                          // src/bsdfs/conductor.cpp:155-177
		d.type = MTSB_BSDF_CONDUCTOR;
		ref<FileResolver> fResolver = Thread::getThread()->getFileResolver();
		const std::string materialName = props.getString("material", "Cu");
		Spectrum intEta, intK;
		if (to_lower_copy(materialName) == "none") { intEta = Spectrum(0.0f); intK = Spectrum(1.0f); }
		else {
			intEta.fromContinuousSpectrum(InterpolatedSpectrum(fResolver->resolve(fs::pathstr("data/ior/" + materialName + ".eta.spd"))));
			intK.fromContinuousSpectrum(InterpolatedSpectrum(fResolver->resolve(fs::pathstr("data/ior/" + materialName + ".k.spd"))));
		}
		const Float extEta = lookupIOR(props, "extEta", "air");
		toRGB(props.getSpectrum("eta", intEta) / extEta, d.p + 0);
		toRGB(props.getSpectrum("k", intK) / extEta, d.p + 3);
		toRGB(props.getSpectrum("specularReflectance", Spectrum(1.0f)), d.p + 6);
	} else if (cls == "TwoSidedBRDF") {                             // src/bsdfs/twosided.cpp:67-186 around any BRDF of the subset
		// The adapter keeps its children in `protected: ref<BSDF> m_nestedBRDF[2]` (twosided.cpp:232), its only data members,
		// and offers no accessor; the class lives in a .cpp, so it cannot be subclassed either.  Its object is therefore read
		// through a layout mirror — BSDF base, then two pointers — compiled with the reference's own headers and flags
		// (plugin/Makefile).  What comes out is checked twice: both must be BSDFs by RTTI, and selfCheckBSDF() compares the resulting device BSDF with TwoSidedBRDF::eval/pdf at load.
		struct Mirror { char base[sizeof(BSDF)]; const BSDF *nested[2]; };
		const Mirror *mir = reinterpret_cast<const Mirror *>(static_cast<const BSDF *>(bsdf));
		const BSDF *n0 = mir->nested[0], *n1 = mir->nested[1];
		const std::string desc = bsdf->toString();
		if (!n0 || !n1 || !n0->getClass()->derivesFrom(MTS_CLASS(BSDF)) || !n1->getClass()->derivesFrom(MTS_CLASS(BSDF)))
			SLog(EError, "gpupath: cannot read the nested BRDFs of %s", desc.substr(0, 200).c_str());
		mtsb_bsdf_desc d1;
		if (!describeBSDF(n0, d, table, texture) || !describeBSDF(n1, d1))
			SLog(EError, "gpupath: twosided around a BRDF outside the GPU subset: %s", desc.c_str());
		if (memcmp(&d, &d1, sizeof(d)) != 0)
			SLog(EError, "gpupath: twosided with two DIFFERENT nested BRDFs is outside the GPU subset: %s", desc.c_str());
		if (d.type & MTSB_BSDF_TWOSIDED_FLAG) SLog(EError, "gpupath: nested twosided adapters are not supported");
		d.type |= MTSB_BSDF_TWOSIDED_FLAG;
	} else {
		return false;
	}
	return true;
}

/// Cross-checks the device BSDF against the CPU plugin on a few fixed directions (SURVEY.md §7 "self-check at load").
static void selfCheckBSDF(mtsb_scene *scene, int id, const BSDF *bsdf) {
	static const float dirs[4][6] = {
		{ 0.3f, 0.2f, 0.93273791f, -0.4f, 0.1f, 0.91104336f }, { 0.0f, 0.0f, 1.0f, 0.6f, 0.0f, 0.8f },
		{ -0.5f, 0.5f, 0.70710678f, 0.5f, -0.5f, 0.70710678f }, { 0.8f, 0.1f, 0.59160798f, -0.7f, -0.2f, 0.68556546f } };
	float in[4][8], out[4][16];
	for (int i = 0; i < 4; ++i) { memcpy(in[i], dirs[i], 6 * sizeof(float)); in[i][6] = 0.35f + 0.1f * i; in[i][7] = 0.6f - 0.1f * i; }
	MTSB_CALL(mtsb_eval_bsdf(scene, id, &in[0][0], 4, &out[0][0]));
	Intersection its = dummyIntersection();
	for (int i = 0; i < 4; ++i) {
		Vector wi(in[i][0], in[i][1], in[i][2]), wo(in[i][3], in[i][4], in[i][5]);
		its.wi = wi;
		BSDFSamplingRecord bRec(its, wi, wo, ERadiance);
		float ref[3]; toRGB(bsdf->eval(bRec), ref);
		const Float pdf = bsdf->pdf(bRec);
		for (int c = 0; c < 3; ++c)
			if (std::abs(ref[c] - out[i][c]) > 2e-3f * (std::abs(ref[c]) + 1e-3f))
				SLog(EError, "gpupath: device BSDF %d (%s) disagrees with the CPU plugin: eval[%d] = %g vs %g", id,
					bsdf->getClass()->getName().c_str(), c, out[i][c], ref[c]);
		if (std::abs(pdf - out[i][3]) > 2e-3f * (std::abs(pdf) + 1e-3f))
			SLog(EError, "gpupath: device BSDF %d (%s) disagrees with the CPU plugin: pdf = %g vs %g", id,
				bsdf->getClass()->getName().c_str(), out[i][3], (float) pdf);
	}
}

MTS_NAMESPACE_END
#include "extract_scene.h"
