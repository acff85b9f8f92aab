// ply.cu — mesh files decoded on the device: binary PLY and `serialized` (the on-disk geometry path).
//
// What the reference does on one host thread (src/shapes/ply.cpp:146-290 driving src/shapes/ply/ply_parser.cpp: one
// std::function call per scalar, per vertex and per face, then TriMesh::configure -> computeNormals,
// src/librender/trimesh.cpp:608-681) becomes three streaming kernels over the file image in HBM:
//
//   ply_vertex_kernel   one thread per vertex record: x y z through toWorld (Transform::operator()(Point), transform.h:108-125),
//                       nx ny nz through the inverse transpose and normalised (ply.cpp:205-212), u v copied
//   ply_face_kernel     one thread per face record: the list count is checked against the layout (3 or 4 everywhere, see
//                       below), indices are bounds-checked, a quad becomes (0 1 2) (3 0 2) (ply.cpp:276-289); when the file
//                       has no normals the angle-weighted face normals are accumulated per vertex (trimesh.cpp:636-658)
//   ply_normalize_kernel   the division of trimesh.cpp:660-672, (1,0,0) for a vertex without a usable normal
//
// The host parses the text header only.  Records are read at arbitrary byte offsets (a 13-byte face record is never
// aligned) as two aligned 32-bit words and a funnel shift; neighbouring threads share those words through L1, so DRAM sees
// each byte of the file once: the kernels are HBM-streaming, algorithmic bytes = file body + 32 B per vertex + 12 B per
// triangle written.
//
// A face list is variable-length in the format, which makes the position of face i depend on every face before it.  The two
// layouts that make up practically every file — all triangles or all quads, told apart by the size of the face element — need
// no index at all (every count byte is still verified on the device); for a file that mixes the two the host makes ONE pass
// over the count fields and hands the kernel a {record offset, first triangle} pair per face.
#include "scene.cuh"
#include <string>
#include <cstring>
#include <cstdlib>
#include <zlib.h>

namespace mtsb {

struct PlyLayout {
	uint32_t vertexCount, faceCount;
	uint32_t vertexStride, faceStride;
	int offPos[3], offNrm[3], offUV[2];   // byte offsets inside a vertex record, -1 = absent
	uint32_t faceListOffset;              // offset of the list inside a face record
	uint32_t faceCountBytes;              // 1 (uchar) or 4 (uint); 0 = no count field (`serialized`: bare index triples)
	uint32_t vertsPerFace;                // 3 or 4
	size_t vertexBegin, faceBegin;        // byte offsets of the two elements in the file
	// PLY interleaves a vertex's properties in one record (the three bases equal vertexBegin, the strides vertexStride); the
	// `serialized` payload stores one array per attribute
	size_t posBase, nrmBase, uvBase;
	uint32_t posStride, nrmStride, uvStride;
	int f64;                              // attributes are doubles (`serialized` with EDoublePrecision), converted like readHelper does
	int normalizeStored;                  // ply.cpp:210 always normalises a stored normal, serialized.cpp:191-194 only behind a toWorld
	int bigEndian;                        // `format binary_big_endian`: every multi-byte field is byte-swapped after the load
	// a file that mixes triangles and quads: per face {byte offset of its record inside the face element, index of its first
	// triangle}, found by ONE sequential pass of the host over the count bytes (records are self-delimiting only in order);
	// NULL for the uniform layouts, where both follow from the face's number
	const uint2 *faceIndex;
	uint32_t triangleCount;
};

__device__ __forceinline__ uint32_t load_u32(const uint32_t *words, size_t byteOffset) {
	const size_t w = byteOffset >> 2; const uint32_t sh = (uint32_t) (byteOffset & 3) * 8;
	const uint32_t lo = words[w];
	if (sh == 0) return lo;
	return __funnelshift_r(lo, words[w + 1], sh);
}
__device__ __forceinline__ uint32_t load_u32(const uint32_t *words, size_t byteOffset, int bigEndian) {
	const uint32_t v = load_u32(words, byteOffset);
	return bigEndian ? __byte_perm(v, 0, 0x0123) : v;
}
__device__ __forceinline__ float load_f32(const uint32_t *words, size_t byteOffset) { return __uint_as_float(load_u32(words, byteOffset)); }
__device__ __forceinline__ float load_real(const uint32_t *words, size_t byteOffset, int f64, int bigEndian = 0) {
	if (!f64) return __uint_as_float(load_u32(words, byteOffset, bigEndian));
	return (float) __hiloint2double((int) load_u32(words, byteOffset + 4), (int) load_u32(words, byteOffset));     // (Float) temp[i]
}

struct PlyXform { float m[16], inv[16]; int identity; };

__global__ void ply_vertex_kernel(const uint32_t *words, size_t base, PlyLayout L, PlyXform X, int flipNormals, int zeroNormals,
		float *positions, float *normals, float *texcoords) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= L.vertexCount) return;
	const size_t recP = base + L.posBase + (size_t) i * L.posStride, recN = base + L.nrmBase + (size_t) i * L.nrmStride,
		recT = base + L.uvBase + (size_t) i * L.uvStride;
	float p[3];
	for (int a = 0; a < 3; ++a) p[a] = L.offPos[a] >= 0 ? load_real(words, recP + L.offPos[a], L.f64, L.bigEndian) : 0.0f;
	if (!X.identity) {
		// the reference's expression order, no contraction
		float q[4];
		for (int r = 0; r < 4; ++r)
			q[r] = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(X.m[r*4], p[0]), __fmul_rn(X.m[r*4+1], p[1])), __fmul_rn(X.m[r*4+2], p[2])), X.m[r*4+3]);
		if (q[3] != 1.0f) { const float inv = 1.0f / q[3]; q[0] *= inv; q[1] *= inv; q[2] *= inv; }   // Point / w multiplies by the reciprocal (point.h)
		p[0] = q[0]; p[1] = q[1]; p[2] = q[2];
	}
	positions[(size_t) i * 3] = p[0]; positions[(size_t) i * 3 + 1] = p[1]; positions[(size_t) i * 3 + 2] = p[2];
	if (L.offNrm[0] >= 0 && !zeroNormals) {
		float n[3];
		for (int a = 0; a < 3; ++a) n[a] = L.offNrm[a] >= 0 ? load_real(words, recN + L.offNrm[a], L.f64, L.bigEndian) : 0.0f;
		if (!X.identity) {
			float t[3];
			for (int c = 0; c < 3; ++c)
				t[c] = __fadd_rn(__fadd_rn(__fmul_rn(X.inv[c], n[0]), __fmul_rn(X.inv[4 + c], n[1])), __fmul_rn(X.inv[8 + c], n[2]));
			n[0] = t[0]; n[1] = t[1]; n[2] = t[2];
		}
		float inv = L.normalizeStored ? 1.0f / sqrtf(n[0]*n[0] + n[1]*n[1] + n[2]*n[2]) : 1.0f;     // normalize(): v / length
		if (flipNormals) inv = -inv;                                      // computeNormals: m_normals[i] *= -1
		normals[(size_t) i * 3] = n[0] * inv; normals[(size_t) i * 3 + 1] = n[1] * inv; normals[(size_t) i * 3 + 2] = n[2] * inv;
	} else {
		normals[(size_t) i * 3] = 0; normals[(size_t) i * 3 + 1] = 0; normals[(size_t) i * 3 + 2] = 0;
	}
	texcoords[(size_t) i * 2]     = L.offUV[0] >= 0 ? load_real(words, recT + L.offUV[0], L.f64, L.bigEndian) : 0.0f;
	texcoords[(size_t) i * 2 + 1] = L.offUV[1] >= 0 ? load_real(words, recT + L.offUV[1], L.f64, L.bigEndian) : 0.0f;
}

__device__ __forceinline__ float unit_angle(const float *u, const float *v) {      // include/mitsuba/core/util.h:321-326
	const float d = u[0]*v[0] + u[1]*v[1] + u[2]*v[2];
	if (d < 0) { const float x = v[0]+u[0], y = v[1]+u[1], z = v[2]+u[2]; return 3.14159265358979323846f - 2.0f * asinf(0.5f * sqrtf(x*x + y*y + z*z)); }
	const float x = v[0]-u[0], y = v[1]-u[1], z = v[2]-u[2];
	return 2.0f * asinf(0.5f * sqrtf(x*x + y*y + z*z));
}

// positions of THIS mesh (already decoded); accumulates into normals of this mesh
__device__ void accumulate_face_normal(const uint32_t ix[3], const float *positions, float *normals) {
	float P[3][3];
	for (int k = 0; k < 3; ++k) for (int a = 0; a < 3; ++a) P[k][a] = positions[(size_t) ix[k] * 3 + a];
	float n[3] = { 0, 0, 0 };
	for (int i = 0; i < 3; ++i) {
		const float *v0 = P[i], *v1 = P[(i + 1) % 3], *v2 = P[(i + 2) % 3];
		float A[3] = { v1[0]-v0[0], v1[1]-v0[1], v1[2]-v0[2] }, B[3] = { v2[0]-v0[0], v2[1]-v0[1], v2[2]-v0[2] };
		if (i == 0) {
			n[0] = A[1]*B[2] - A[2]*B[1]; n[1] = A[2]*B[0] - A[0]*B[2]; n[2] = A[0]*B[1] - A[1]*B[0];
			const float len = sqrtf(n[0]*n[0] + n[1]*n[1] + n[2]*n[2]);
			if (len == 0) break;
			const float inv = 1.0f / len; n[0] *= inv; n[1] *= inv; n[2] *= inv;
		}
		const float la = 1.0f / sqrtf(A[0]*A[0] + A[1]*A[1] + A[2]*A[2]), lb = 1.0f / sqrtf(B[0]*B[0] + B[1]*B[1] + B[2]*B[2]);
		for (int a = 0; a < 3; ++a) { A[a] *= la; B[a] *= lb; }
		const float angle = unit_angle(A, B);
		float *dst = normals + (size_t) ix[i] * 3;
		atomicAdd(dst, n[0] * angle); atomicAdd(dst + 1, n[1] * angle); atomicAdd(dst + 2, n[2] * angle);
	}
}

// error word: 1 = a face's list count differs from the layout, 2 = an index is out of range
__global__ void ply_face_kernel(const uint32_t *words, size_t base, PlyLayout L, int swapWinding, int accumulate,
		const float *positions, float *normals, uint32_t *indices, uint32_t *error) {
	const uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
	if (f >= L.faceCount) return;
	size_t rec = base + L.faceBegin + (size_t) f * L.faceStride + L.faceListOffset;
	uint32_t vertsPerFace = L.vertsPerFace, firstTri = f * (L.vertsPerFace - 2);
	if (L.faceIndex) { const uint2 fi = L.faceIndex[f]; rec = base + L.faceBegin + fi.x + L.faceListOffset; firstTri = fi.y; }
	if (L.faceCountBytes) {
		uint32_t count = load_u32(words, rec);                 // a one-byte count is the first byte in memory either way
		if (L.faceCountBytes == 1) count &= 0xffu;
		else if (L.bigEndian) count = __byte_perm(count, 0, 0x0123);
		if (L.faceIndex) { if (count != 3 && count != 4) { atomicOr(error, 1u); return; } vertsPerFace = count; }
		else if (count != L.vertsPerFace) { atomicOr(error, 1u); return; }
	}
	uint32_t v[4] = { 0, 0, 0, 0 };
	bool ok = true;
	for (uint32_t k = 0; k < vertsPerFace; ++k) {
		v[k] = load_u32(words, rec + L.faceCountBytes + 4 * k, L.bigEndian);
		ok = ok && v[k] < L.vertexCount;          // Assert((size_t) element < m_vertexCount), ply.cpp:266,272 (a negative int32 fails it too)
	}
	if (!ok) { atomicOr(error, 2u); return; }
	const uint32_t trisPerFace = vertsPerFace - 2;
	uint32_t *out = indices + (size_t) firstTri * 3;
	uint32_t t0[3] = { v[0], v[1], v[2] }, t1[3] = { v[3], v[0], v[2] };
	if (swapWinding) { uint32_t s = t0[0]; t0[0] = t0[1]; t0[1] = s; s = t1[0]; t1[0] = t1[1]; t1[1] = s; }   // faceNormals + flipNormals (trimesh.cpp:616-622)
	out[0] = t0[0]; out[1] = t0[1]; out[2] = t0[2];
	if (accumulate) accumulate_face_normal(t0, positions, normals);
	if (trisPerFace == 2) {
		out[3] = t1[0]; out[4] = t1[1]; out[5] = t1[2];
		if (accumulate) accumulate_face_normal(t1, positions, normals);
	}
}

__global__ void ply_normalize_kernel(uint32_t vertexCount, int flipNormals, float *normals, uint32_t *invalid) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= vertexCount) return;
	float *n = normals + (size_t) i * 3;
	float len = sqrtf(n[0]*n[0] + n[1]*n[1] + n[2]*n[2]);
	if (flipNormals) len = -len;
	if (len != 0) { const float inv = 1.0f / len; n[0] *= inv; n[1] *= inv; n[2] *= inv; }
	else { n[0] = 1; n[1] = 0; n[2] = 0; atomicAdd(invalid, 1u); }
}

// ---------------------------------------------------------------------------------------------------------------------------
// host: header
// ---------------------------------------------------------------------------------------------------------------------------
static int scalar_size(const std::string &t) {
	if (t == "char" || t == "int8" || t == "uchar" || t == "uint8") return 1;
	if (t == "short" || t == "int16" || t == "ushort" || t == "uint16") return 2;
	if (t == "int" || t == "int32" || t == "uint" || t == "uint32" || t == "float" || t == "float32") return 4;
	if (t == "double" || t == "float64") return 8;
	return 0;
}
static bool is_float32(const std::string &t) { return t == "float" || t == "float32"; }

static std::vector<std::string> split_ws(const std::string &line) {
	std::vector<std::string> out; size_t i = 0;
	while (i < line.size()) {
		while (i < line.size() && (line[i] == ' ' || line[i] == '\t' || line[i] == '\r')) ++i;
		size_t j = i;
		while (j < line.size() && line[j] != ' ' && line[j] != '\t' && line[j] != '\r') ++j;
		if (j > i) out.push_back(line.substr(i, j - i));
		i = j;
	}
	return out;
}

struct PlyMesh {
	PlyLayout L; PlyXform X;
	DevBuf<uint32_t> raw; size_t base = 0;       // file image (PLY) / inflated payload (serialized) on the device, its first byte at byte `base` of raw
	int shape = -1; bool hasNormals = false, faceNormals = false, flipNormals = false;
	bool mirrored = false;                       // serialized.cpp:197-202: a toWorld with a negative determinant swaps the winding
	const char *kind = "PLY";
	DevBuf<uint2> faceIndex;                     // mixed triangle / quad files only
	uint32_t vtxBase = 0, triBase = 0;
};

static int parse_header(const unsigned char *file, size_t bytes, PlyLayout &L, bool &hasNormals, bool &hasUV, std::vector<uint2> *faceIndexOut = nullptr) {
	memset(&L, 0, sizeof(L));
	for (int a = 0; a < 3; ++a) { L.offPos[a] = -1; L.offNrm[a] = -1; }
	L.offUV[0] = L.offUV[1] = -1;
	size_t pos = 0; int lineNo = 0;
	enum { NONE, VERTEX, FACE, OTHER } cur = NONE;
	bool bigEndian = false, sawFormat = false, sawVertex = false, sawFace = false, faceList = false, done = false;
	size_t otherBytesAfterFace = 0, otherCount = 0, otherStride = 0; bool otherBeforeVertex = false;
	size_t bytesBeforeVertex = 0, bytesBetween = 0;
	auto closeOther = [&]() {
		if (cur != OTHER) return;
		const size_t b = otherCount * otherStride;
		if (!sawVertex) bytesBeforeVertex += b; else if (!sawFace) bytesBetween += b; else otherBytesAfterFace += b;
		(void) otherBeforeVertex;
	};
	while (pos < bytes && !done) {
		size_t e = pos; while (e < bytes && file[e] != '\n') ++e;
		if (e >= bytes) break;
		std::string line((const char *) file + pos, e - pos); pos = e + 1; ++lineNo;
		std::vector<std::string> w = split_ws(line);
		if (lineNo == 1) { if (w.size() != 1 || w[0] != "ply") MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_ply: not a PLY file (no 'ply' magic)"); continue; }
		if (w.empty() || w[0] == "comment" || w[0] == "obj_info") continue;
		if (w[0] == "format") {
			if (w.size() < 3 || (w[1] != "binary_little_endian" && w[1] != "binary_big_endian"))
				MTSB_FAIL(MTSB_EUNSUPPORTED, "mtsb_scene_add_ply: format '%s' (the device decoder reads binary_little_endian and binary_big_endian; ascii files go through the host parser)", w.size() > 1 ? w[1].c_str() : "");
			sawFormat = true; bigEndian = w[1] == "binary_big_endian";
		} else if (w[0] == "element") {
			if (w.size() != 3) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_ply: header line %d: malformed element", lineNo);
			closeOther();
			const size_t count = (size_t) strtoull(w[2].c_str(), nullptr, 10);
			if (w[1] == "vertex") {
				if (sawVertex || sawFace) MTSB_FAIL(MTSB_EUNSUPPORTED, "mtsb_scene_add_ply: 'vertex' element after 'face' or repeated");
				if (count > 0xffffffffull) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_ply: too many vertices");
				cur = VERTEX; sawVertex = true; L.vertexCount = (uint32_t) count;
			} else if (w[1] == "face") {
				if (!sawVertex || sawFace) MTSB_FAIL(MTSB_EUNSUPPORTED, "mtsb_scene_add_ply: 'face' element before 'vertex' or repeated");
				if (count > 0x7fffffffull) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_ply: too many faces");
				cur = FACE; sawFace = true; L.faceCount = (uint32_t) count;
			} else { cur = OTHER; otherCount = count; otherStride = 0; }
		} else if (w[0] == "property") {
			if (cur == NONE || w.size() < 3) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_ply: header line %d: property outside an element", lineNo);
			if (w[1] == "list") {
				if (w.size() != 5) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_ply: header line %d: malformed list property", lineNo);
				if (cur != FACE || faceList || (w[4] != "vertex_indices" && w[4] != "vertex_index"))
					MTSB_FAIL(MTSB_EUNSUPPORTED, "mtsb_scene_add_ply: list property '%s' (only the face element's vertex_indices list has a fixed record size)", w[4].c_str());
				const int cs = scalar_size(w[2]), is = scalar_size(w[3]);
				// the reference binds the list for <uint8|uint32 size, int32|uint32 index> (ply.cpp:366-440)
				if (!((cs == 1 && (w[2] == "uchar" || w[2] == "uint8")) || (cs == 4 && (w[2] == "uint" || w[2] == "uint32"))) ||
				    !(is == 4 && (w[3] == "int" || w[3] == "int32" || w[3] == "uint" || w[3] == "uint32")))
					MTSB_FAIL(MTSB_EUNSUPPORTED, "mtsb_scene_add_ply: vertex_indices as 'list %s %s'", w[2].c_str(), w[3].c_str());
				faceList = true; L.faceListOffset = L.faceStride; L.faceCountBytes = (uint32_t) cs;
				// the list's own bytes are added once the face count per record is known
			} else {
				const int sz = scalar_size(w[1]);
				if (!sz) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_ply: header line %d: unknown type '%s'", lineNo, w[1].c_str());
				if (cur == VERTEX) {
					const std::string &n = w[2]; const int off = (int) L.vertexStride;
					int *slot = nullptr;
					if (n == "x") slot = &L.offPos[0]; else if (n == "y") slot = &L.offPos[1]; else if (n == "z") slot = &L.offPos[2];
					else if (n == "nx") slot = &L.offNrm[0]; else if (n == "ny") slot = &L.offNrm[1]; else if (n == "nz") slot = &L.offNrm[2];
					else if (n == "u" || n == "texture_u" || n == "s") slot = &L.offUV[0];
					else if (n == "v" || n == "texture_v" || n == "t") slot = &L.offUV[1];
					// like the reference, a geometry property that is not float32 has no callback and stays at its default
					if (slot && is_float32(w[1])) *slot = off;
					L.vertexStride += sz;
				} else if (cur == FACE) {
					if (faceList) MTSB_FAIL(MTSB_EUNSUPPORTED, "mtsb_scene_add_ply: scalar face property after the index list");
					L.faceStride += sz;
				} else otherStride += sz;
			}
		} else if (w[0] == "end_header") {
			closeOther(); done = true;
		} else MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_ply: header line %d: unknown keyword '%s'", lineNo, w[0].c_str());
	}
	if (!done) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_ply: no end_header");
	if (!sawFormat || !sawVertex || !sawFace || !faceList) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_ply: header lacks format, vertex element or face index list");
	if (L.vertexCount == 0 || L.faceCount == 0) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_ply: no triangles or vertices found");   // ply.cpp:101-102
	// nx without the callback that allocates m_normals (normal_x_callback) leaves the mesh without normals: all three or none here
	hasNormals = L.offNrm[0] >= 0;
	hasUV = L.offUV[0] >= 0;       // texcoord_u_callback allocates
	L.vertexBegin = pos + bytesBeforeVertex;
	L.faceBegin = L.vertexBegin + (size_t) L.vertexCount * L.vertexStride + bytesBetween;
	if (L.faceBegin + otherBytesAfterFace > bytes) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_ply: file is shorter than its header says");
	const size_t faceBytes = bytes - L.faceBegin - otherBytesAfterFace;
	const size_t fixed = (size_t) L.faceStride + L.faceCountBytes;
	L.faceIndex = nullptr;
	if (faceBytes == (fixed + 12) * L.faceCount) L.vertsPerFace = 3;
	else if (faceBytes == (fixed + 16) * L.faceCount) L.vertsPerFace = 4;
	else {
		// neither all triangles nor all quads: one pass over the count fields finds where every record starts
		if (faceBytes > 0xffffffffull) MTSB_FAIL(MTSB_EUNSUPPORTED, "mtsb_scene_add_ply: a mixed triangle / quad face element larger than 4 GB");
		const unsigned char *fb = file + L.faceBegin;
		size_t off = 0; uint64_t tri = 0;
		if (faceIndexOut) faceIndexOut->resize(L.faceCount);
		for (uint32_t f = 0; f < L.faceCount; ++f) {
			if (off + fixed > faceBytes) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_ply: %u faces in %zu bytes: the face element ends inside face %u (a truncated file?)", L.faceCount, faceBytes, f);
			const unsigned char *c = fb + off + L.faceListOffset;
			uint32_t count = c[0];
			if (L.faceCountBytes == 4) count = bigEndian ? ((uint32_t) c[0] << 24 | (uint32_t) c[1] << 16 | (uint32_t) c[2] << 8 | c[3]) : ((uint32_t) c[3] << 24 | (uint32_t) c[2] << 16 | (uint32_t) c[1] << 8 | c[0]);
			if (count != 3 && count != 4)      // ply.cpp:251-262
				MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_ply: %u faces in %zu bytes: neither all triangles nor all quads, and face %u has %u vertices (only triangle and quad-based PLY meshes are supported; a truncated file?)", L.faceCount, faceBytes, f, count);
			if (faceIndexOut) (*faceIndexOut)[f] = make_uint2((uint32_t) off, (uint32_t) tri);
			off += fixed + 4 * (size_t) count; tri += count - 2;
		}
		if (off != faceBytes) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_ply: neither all triangles nor all quads, and the faces take %zu bytes of a %zu-byte face element (a truncated file?)", off, faceBytes);
		if (tri > 0x7fffffffull) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_ply: too many triangles");
		L.vertsPerFace = 0; L.triangleCount = (uint32_t) tri;
	}
	if (L.vertsPerFace) { L.faceStride = (uint32_t) (fixed + 4 * L.vertsPerFace); L.triangleCount = L.faceCount * (L.vertsPerFace - 2); }
	else L.faceStride = 0;
	L.posBase = L.nrmBase = L.uvBase = L.vertexBegin;
	L.posStride = L.nrmStride = L.uvStride = L.vertexStride;
	L.f64 = 0; L.normalizeStored = 1; L.bigEndian = bigEndian ? 1 : 0;
	return 0;
}

static int set_transform(PlyMesh *M, const float *toWorld, const float *toWorldInv, const char *who) {
	memset(&M->X, 0, sizeof(M->X));
	M->X.identity = 1;
	for (int i = 0; i < 16; ++i) { M->X.m[i] = M->X.inv[i] = (i % 5 == 0) ? 1.0f : 0.0f; }
	if (toWorld) {
		if (!toWorldInv) MTSB_FAIL(MTSB_EINVAL, "%s: to_world needs its inverse beside it (Transform holds both)", who);
		for (int i = 0; i < 16; ++i) { M->X.m[i] = toWorld[i]; M->X.inv[i] = toWorldInv[i]; if (toWorld[i] != ((i % 5 == 0) ? 1.0f : 0.0f)) M->X.identity = 0; }
	}
	return 0;
}

// the image goes to the device with its position block on a 16-byte boundary; 8 bytes of slack behind it for load_u32
static int upload_image(mtsb_scene *scene, PlyMesh *M, const void *image, size_t bytes, const char *who) {
	mtsb_ctx *ctx = scene->ctx;
	cudaSetDevice(ctx->device);
	M->base = (16 - (M->L.posBase & 15)) & 15;
	const size_t words = (M->base + bytes + 8 + 3) / 4;
	MTSB_CHECK(M->raw.alloc(words));
	cudaError_t e = cudaMemsetAsync(M->raw.ptr + (words - 3), 0, 12, ctx->stream);
	if (e == cudaSuccess) e = cudaMemcpyAsync((unsigned char *) M->raw.ptr + M->base, image, bytes, cudaMemcpyHostToDevice, ctx->stream);
	if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);          // the caller's buffer may go away after the call
	if (e != cudaSuccess) MTSB_FAIL(MTSB_ECUDA, "%s: upload: %s", who, cudaGetErrorString(e));
	return 0;
}

int ply_stage(mtsb_scene *scene, const void *file, size_t bytes, const float *toWorld, const float *toWorldInv, uint32_t flags, int bsdf,
		uint32_t *vertexCountOut, uint32_t *triangleCountOut, bool *hasUVOut, bool *hasNormalsOut, PlyMesh **out) {
	PlyMesh *M = new PlyMesh();
	bool hasUV = false;
	std::vector<uint2> faceIndex;
	int rc = parse_header((const unsigned char *) file, bytes, M->L, M->hasNormals, hasUV, &faceIndex);
	if (!rc) rc = set_transform(M, toWorld, toWorldInv, "mtsb_scene_add_ply");
	M->faceNormals = (flags & MTSB_PLY_FACE_NORMALS) != 0; M->flipNormals = (flags & MTSB_PLY_FLIP_NORMALS) != 0;
	if (!rc && M->L.vertsPerFace == 0) {
		cudaSetDevice(scene->ctx->device);
		rc = M->faceIndex.upload(faceIndex.data(), faceIndex.size(), scene->ctx->stream);
		M->L.faceIndex = M->faceIndex.ptr;
	}
	if (!rc) rc = upload_image(scene, M, file, bytes, "mtsb_scene_add_ply");      // synchronizes: faceIndex may go out of scope
	if (rc) { delete M; return rc; }
	(void) bsdf;
	*vertexCountOut = M->L.vertexCount; *triangleCountOut = M->L.triangleCount;
	*hasUVOut = hasUV; *hasNormalsOut = M->hasNormals && !M->faceNormals;
	*out = M;
	return 0;
}

// ---------------------------------------------------------------------------------------------------------------------------
// `serialized` (TriMesh::loadCompressed, src/librender/trimesh.cpp:176-253, readOffset :274-296; shapes/serialized.cpp:148-210)
// A deflate stream is sequential: it is inflated here on the host (zlib, as the reference's ZStream does) straight into one
// buffer; the payload — one array per attribute, float or double — goes to the device in one copy and the kernels above
// convert, transform and triangulate it.
// ---------------------------------------------------------------------------------------------------------------------------
static inline uint32_t rd_u32(const unsigned char *p) { uint32_t v; memcpy(&v, p, 4); return v; }
static inline uint64_t rd_u64(const unsigned char *p) { uint64_t v; memcpy(&v, p, 8); return v; }
static inline uint16_t rd_u16(const unsigned char *p) { uint16_t v; memcpy(&v, p, 2); return v; }

static int ser_open(const unsigned char *f, size_t bytes, int index, size_t &dataBegin, int &version) {
	if (bytes < 8) MTSB_FAIL(MTSB_EINVAL, "serialized: file too short");
	if (rd_u16(f) != 0x041C) MTSB_FAIL(MTSB_EINVAL, "serialized: encountered an invalid file format");
	version = rd_u16(f + 2);
	if (version != 3 && version != 4) MTSB_FAIL(MTSB_EINVAL, "serialized: encountered an incompatible file version (%d)", version);
	size_t offset = 0;
	if (index != 0) {
		const uint32_t count = rd_u32(f + bytes - 4);
		if (index < 0 || (uint32_t) index >= count) MTSB_FAIL(MTSB_EINVAL, "serialized: shape index is out of range (requested %d out of 0..%u)", index, count - 1);
		const size_t entry = version == 4 ? 8 : 4;
		if ((size_t) count * entry + 4 > bytes) MTSB_FAIL(MTSB_EINVAL, "serialized: offset dictionary larger than the file");
		const size_t pos = version == 4 ? bytes - 8 * (size_t) (count - index) - 4 : bytes - 4 * (size_t) (count - index + 1);
		offset = version == 4 ? (size_t) rd_u64(f + pos) : (size_t) rd_u32(f + pos);
		if (offset + 4 > bytes) MTSB_FAIL(MTSB_EINVAL, "serialized: mesh offset beyond the end of the file");
	}
	dataBegin = offset + 4;
	return 0;
}

// inflates the mesh's header (flags, name, counts), then — unless probing — the whole payload into `out`
static int ser_inflate(const unsigned char *src, size_t srcBytes, int version, PlyLayout &L, bool &hasNormals, bool &hasUV,
		std::vector<unsigned char> *out) {
	z_stream z; memset(&z, 0, sizeof(z));
	if (inflateInit2(&z, 15) != Z_OK) MTSB_FAIL(MTSB_EINVAL, "serialized: could not initialize zlib");
	z.next_in = (Bytef *) src; z.avail_in = (uInt) std::min<size_t>(srcBytes, 0x40000000u);
	size_t consumed = z.avail_in;
	std::vector<unsigned char> head(4096);
	z.next_out = head.data(); z.avail_out = (uInt) head.size();
	int rc = inflate(&z, Z_SYNC_FLUSH);
	if (rc != Z_OK && rc != Z_STREAM_END) { inflateEnd(&z); MTSB_FAIL(MTSB_EINVAL, "serialized: inflate(): stream error %d", rc); }
	const size_t got = head.size() - z.avail_out;
	size_t p = 4;
	if (version == 4) { while (p < got && head[p] != 0) ++p; ++p; }
	if (p + 16 > got) { inflateEnd(&z); MTSB_FAIL(MTSB_EINVAL, "serialized: mesh header truncated (or a name longer than 4 KB)"); }
	const uint32_t flags = rd_u32(head.data());
	const uint64_t V = rd_u64(head.data() + p), T = rd_u64(head.data() + p + 8);
	p += 16;
	if (V == 0 || T == 0 || V > 0xffffffffull || T > 0x7fffffffull) { inflateEnd(&z); MTSB_FAIL(MTSB_EINVAL, "serialized: %llu vertices, %llu triangles", (unsigned long long) V, (unsigned long long) T); }
	memset(&L, 0, sizeof(L));
	L.f64 = (flags & 0x2000u) ? 1 : 0;                                    // EDoublePrecision
	const uint32_t fsz = L.f64 ? 8 : 4;
	L.vertexCount = (uint32_t) V; L.faceCount = (uint32_t) T; L.triangleCount = (uint32_t) T; L.faceIndex = nullptr; L.vertsPerFace = 3; L.faceCountBytes = 0; L.faceListOffset = 0; L.faceStride = 12;
	for (int a = 0; a < 3; ++a) { L.offPos[a] = (int) (a * fsz); L.offNrm[a] = -1; }
	L.offUV[0] = L.offUV[1] = -1;
	size_t next = p;
	L.posBase = next; L.posStride = 3 * fsz; next += (size_t) V * 3 * fsz;
	L.nrmBase = L.uvBase = L.posBase; L.nrmStride = L.uvStride = 0;
	hasNormals = (flags & 0x0001u) != 0; hasUV = (flags & 0x0002u) != 0;
	if (hasNormals) { L.nrmBase = next; L.nrmStride = 3 * fsz; for (int a = 0; a < 3; ++a) L.offNrm[a] = (int) (a * fsz); next += (size_t) V * 3 * fsz; }
	if (hasUV) { L.uvBase = next; L.uvStride = 2 * fsz; L.offUV[0] = 0; L.offUV[1] = (int) fsz; next += (size_t) V * 2 * fsz; }
	if (flags & 0x0008u) next += (size_t) V * 3 * fsz;                    // EHasColors: stepped over
	L.faceBegin = next; L.vertexBegin = L.posBase; L.vertexStride = L.posStride;
	const size_t total = next + (size_t) T * 12;
	if (!out) { inflateEnd(&z); return 0; }
	// deflate expands by at most 1032 : 1 — counts that promise more than the stream can hold are refused before allocating
	if (total > srcBytes * 1100 + 4096) { inflateEnd(&z); MTSB_FAIL(MTSB_EINVAL, "serialized: header promises %zu payload bytes, the stream has %zu", total, srcBytes); }
	try { out->resize(total); } catch (const std::exception &) { inflateEnd(&z); MTSB_FAIL(MTSB_ENOMEM, "serialized: %zu bytes of host memory for the payload", total); }
	memcpy(out->data(), head.data(), std::min(got, total));
	size_t have = std::min(got, total);
	while (have < total && rc != Z_STREAM_END) {
		if (z.avail_in == 0 && consumed < srcBytes) { z.avail_in = (uInt) std::min<size_t>(srcBytes - consumed, 0x40000000u); consumed += z.avail_in; }
		z.next_out = out->data() + have; z.avail_out = (uInt) std::min<size_t>(total - have, 0x40000000u);
		const size_t before = z.avail_out;
		rc = inflate(&z, Z_NO_FLUSH);
		have += before - z.avail_out;
		if (rc != Z_OK && rc != Z_STREAM_END) break;
		if (before == z.avail_out && z.avail_in == 0 && consumed >= srcBytes) break;
	}
	inflateEnd(&z);
	if (have < total) MTSB_FAIL(MTSB_EINVAL, "serialized: the stream ends after %zu of %zu payload bytes (zlib %d)", have, total, rc);
	return 0;
}

int ser_probe(const void *file, size_t bytes, int index, uint32_t *nv, uint32_t *nt) {
	size_t begin = 0; int version = 0; PlyLayout L; bool hn, hu;
	MTSB_CHECK(ser_open((const unsigned char *) file, bytes, index, begin, version));
	MTSB_CHECK(ser_inflate((const unsigned char *) file + begin, bytes - begin, version, L, hn, hu, nullptr));
	if (nv) *nv = L.vertexCount;
	if (nt) *nt = L.faceCount;
	return 0;
}

int ser_stage(mtsb_scene *scene, const void *file, size_t bytes, int index, const float *toWorld, const float *toWorldInv, uint32_t flags,
		uint32_t *vertexCountOut, uint32_t *triangleCountOut, bool *hasUVOut, bool *hasNormalsOut, PlyMesh **out) {
	size_t begin = 0; int version = 0;
	MTSB_CHECK(ser_open((const unsigned char *) file, bytes, index, begin, version));
	PlyMesh *M = new PlyMesh();
	M->kind = "serialized";
	bool hasUV = false;
	std::vector<unsigned char> payload;
	int rc = ser_inflate((const unsigned char *) file + begin, bytes - begin, version, M->L, M->hasNormals, hasUV, &payload);
	if (!rc) rc = set_transform(M, toWorld, toWorldInv, "mtsb_scene_add_serialized");
	if (!rc) {
		const float *m = M->X.m;
		const float det = m[0] * (m[5] * m[10] - m[6] * m[9]) - m[1] * (m[4] * m[10] - m[6] * m[8]) + m[2] * (m[4] * m[9] - m[5] * m[8]);
		M->mirrored = det < 0;
		M->L.normalizeStored = M->X.identity ? 0 : 1;
		M->faceNormals = (flags & MTSB_PLY_FACE_NORMALS) != 0; M->flipNormals = (flags & MTSB_PLY_FLIP_NORMALS) != 0;
		rc = upload_image(scene, M, payload.data(), payload.size(), "mtsb_scene_add_serialized");
	}
	if (rc) { delete M; return rc; }
	*vertexCountOut = M->L.vertexCount; *triangleCountOut = M->L.faceCount;
	*hasUVOut = hasUV; *hasNormalsOut = M->hasNormals && !M->faceNormals;
	*out = M;
	return 0;
}

int ply_probe(const void *file, size_t bytes, uint32_t *nv, uint32_t *nt) {
	PlyLayout L; bool hn = false, hu = false;
	MTSB_CHECK(parse_header((const unsigned char *) file, bytes, L, hn, hu));
	if (nv) *nv = L.vertexCount;
	if (nt) *nt = L.triangleCount;
	return 0;
}

void ply_release(PlyMesh *M) { delete M; }
void ply_place(PlyMesh *M, int shape, uint32_t vtxBase, uint32_t triBase) { M->shape = shape; M->vtxBase = vtxBase; M->triBase = triBase; }

// commit: decode straight into the scene's vertex / index arrays
int ply_decode(mtsb_scene *scene, PlyMesh *M, float *positions, float *normals, float *texcoords, uint32_t *indices, uint32_t *scratch) {
	cudaStream_t st = scene->ctx->stream;
	const PlyLayout &L = M->L;
	float *P = positions + (size_t) M->vtxBase * 3, *N = normals + (size_t) M->vtxBase * 3, *T = texcoords + (size_t) M->vtxBase * 2;
	uint32_t *I = indices + (size_t) M->triBase * 3;
	const bool compute = !M->faceNormals && !M->hasNormals;          // trimesh.cpp:631-672
	MTSB_CUDA(cudaMemsetAsync(scratch, 0, 2 * sizeof(uint32_t), st));
	cudaEvent_t ev0, ev1;
	MTSB_CUDA(cudaEventCreate(&ev0)); MTSB_CUDA(cudaEventCreate(&ev1));
	MTSB_CUDA(cudaEventRecord(ev0, st));
	ply_vertex_kernel<<<(L.vertexCount + 255) / 256, 256, 0, st>>>(M->raw.ptr, M->base, L, M->X, M->flipNormals ? 1 : 0, (M->faceNormals || compute) ? 1 : 0, P, N, T);
	ply_face_kernel<<<(L.faceCount + 255) / 256, 256, 0, st>>>(M->raw.ptr, M->base, L, ((M->faceNormals && M->flipNormals) != M->mirrored) ? 1 : 0, compute ? 1 : 0, P, N, I, scratch);
	if (compute) ply_normalize_kernel<<<(L.vertexCount + 255) / 256, 256, 0, st>>>(L.vertexCount, M->flipNormals ? 1 : 0, N, scratch + 1);
	MTSB_CUDA(cudaEventRecord(ev1, st));
	MTSB_CUDA(cudaGetLastError());
	uint32_t err[2] = { 0, 0 };
	MTSB_CUDA(cudaMemcpyAsync(err, scratch, sizeof(err), cudaMemcpyDeviceToHost, st));
	MTSB_CUDA(cudaStreamSynchronize(st));
	float ms = 0; cudaEventElapsedTime(&ms, ev0, ev1); scene->plyDecodeMs += ms;
	cudaEventDestroy(ev0); cudaEventDestroy(ev1);
	if (err[0] & 1u) MTSB_FAIL(MTSB_EUNSUPPORTED, "mtsb_scene_commit: %s shape %d mixes triangles and quads (or holds a face that is neither): load it with the host parser", M->kind, M->shape);
	if (err[0] & 2u) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_commit: %s shape %d: a face references a vertex >= %u", M->kind, M->shape, L.vertexCount);
	M->raw.release(); M->faceIndex.release();
	return 0;
}

} // namespace mtsb
