/*
 * hrt.h — C ABI of the B200-native path-tracing core that replaces the CPU render loop of
 * Todegal/HobbyRaytracer.
 *
 * The reference has no FFI / plugin interface: its seam is the free function
 *   static void render(int nThreads, shared_ptr<Texture> background, shared_ptr<Hittable> world,
 *                      const Camera& camera, shared_ptr<Film>& film)        (hobbyraytracer/main.cpp:81-140)
 * called once from main() (main.cpp:176), fed by objects that Scene::loadScene constructs
 * (scene.cpp:127-374).  All geometry/material fields of those objects are private, so the scene crosses
 * this boundary as a *builder*: every hrt_scene_add_* call below corresponds to one std::make_shared<X>(...)
 * construction site of the reference and takes exactly the constructor arguments of X.
 *
 * Conventions: plain pointers and sizes only; every int-returning function returns >= 0 on success
 * (an id where documented) and < 0 on error, with the message available from hrt_last_error().
 * Nothing throws across the ABI.  Host pointers are only read during the call that takes them.
 * The library is CUDA-only: there is no CPU fallback; without a usable sm_100 device every compute
 * entry point fails with HRT_ERR_CUDA.
 *
 * The same builder signatures (with the prefix orc_ / ref_ instead of hrt_) are implemented by the CPU
 * oracle (oracle/oracle.cpp) and by the harness that drives the unmodified reference (oracle/ref_harness.cpp),
 * so parity tests feed the identical scene to all three.
 */
#ifndef HRT_H
#define HRT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HRT_ABI_VERSION 7

/* ---- error codes --------------------------------------------------------------------------- */
#define HRT_OK            0
#define HRT_ERR_ARG      -1   /* bad argument / id out of range                                   */
#define HRT_ERR_STATE    -2   /* call not valid in this state (e.g. render before commit)          */
#define HRT_ERR_CUDA     -3   /* CUDA runtime failure (no device, OOM, launch error)               */
#define HRT_ERR_IO       -4   /* file could not be read / written / parsed                         */

#define HRT_MISS 0xFFFFFFFFu  /* primitive id reported for a ray that hit nothing                  */

/* ---- material kinds: Material subclasses of material.h ------------------------------------- */
enum {
    HRT_MAT_LAMBERTIAN    = 0,  /* Lambertian(albedo)                         material.h:132-157 */
    HRT_MAT_METAL         = 1,  /* Metal(albedo, roughness=p0)                material.h:159-182 */
    HRT_MAT_DIELECTRIC    = 2,  /* Dielectric(ior=p0, roughness=p1)           material.h:199-242 */
    HRT_MAT_ISOTROPIC     = 3,  /* Isotropic(albedo)                          material.h:73-89   */
    HRT_MAT_DIFFUSE_LIGHT = 4,  /* DiffuseLight(colour=albedo, strength=p0)   material.h:91-109  */
    HRT_MAT_PBR           = 5,  /* PBR(albedo, metalness=p1, roughness=p0)    material.cpp:4-28  */
    HRT_MAT_UVTEST        = 6   /* UVTest()                                   material.h:111-130 */
};

/* MatVec3 / MatScalar (material.h:10-58): a constant, or a texture when the *_tex id is >= 0.   */
typedef struct hrt_material {
    int32_t kind;
    float   albedo[3];
    int32_t albedo_tex;   /* -1 = use albedo[]                                                    */
    float   p0;
    int32_t p0_tex;       /* -1 = use p0; scalar-from-texture = length(rgb) (material.h:49-51)    */
    float   p1;
    int32_t p1_tex;
} hrt_material;

/* ---- instance wrappers: translate.h / scale.h / rotateQuat.h / rotateY.h ------------------- */
enum {
    HRT_XF_TRANSLATE   = 0,   /* v = offset xyz                        translate.cpp:7-19          */
    HRT_XF_SCALE       = 1,   /* v = factor xyz                        scale.cpp:11-27             */
    HRT_XF_ROTATE_QUAT = 2,   /* v = quaternion (w, x, y, z)           rotateQuat.cpp:44-66        */
    HRT_XF_ROTATE_Y    = 3,   /* v = (sinTheta, cosTheta, -, -)        rotateY.cpp:44-75           */
    HRT_XF_ROTATE_EULER_DEG = 4, /* v = Euler degrees xyz; converted on the host exactly like
                                    glm::quat(glm::radians(angles)) (scene.cpp:340) into ROTATE_QUAT */
    HRT_XF_ROTATE_Y_DEG = 5   /* v = (degrees); converted like RotateY's ctor (rotateY.cpp:7-9)    */
};
typedef struct hrt_xform {
    int32_t kind;
    float   v[4];
} hrt_xform;

/* Camera(lookFrom, lookAt, up, vfov, aspectRatio, aperture, focusDistance)   camera.h:9-30         */
typedef struct hrt_camera {
    float position[3];
    float look_at[3];
    float up[3];
    float vfov_deg;
    float aspect;          /* Film::getAspectRatio() = float(w)/float(h)   film.h:20-22             */
    float aperture;        /* parsed but unused by the reference (camera.h:34-35)                   */
    float focal_distance;
} hrt_camera;

/* BVH flavours for hrt_scene_commit */
enum {
    HRT_BVH_REFERENCE = 0,  /* bvh.cpp:6-61 topology (median split on a pseudo-random axis, leaves of 1-2),
                               flat world list, fixed left-then-right order: reproduces the reference's
                               visiting order, hence its tie-breaking, exactly                      */
    HRT_BVH_SAH       = 1,  /* binned-SAH BVH over each mesh + top-level BVH over world objects, built on host */
    HRT_BVH_LBVH_GPU  = 2,  /* the whole per-mesh build on the GPU: triangle packing, the bvh.cpp:6-61 topology (one
                               stable radix sort per level; equal keys keep their order where std::sort's introsort
                               permutes them - identical whenever keys differ), leaf records and a Morton-code LBVH
                               over them in the SAH node format; top-level BVH on the host                  */
    HRT_BVH_LBVH_HOST = 3   /* host mirror of HRT_BVH_LBVH_GPU: same rules, arrays bit-identical (validation) */
};

typedef struct hrt_render_params {
    int32_t  width, height;
    int32_t  spp;            /* film.samples: divisor of the per-pixel mean (main.cpp:126)          */
    int32_t  sample_first;   /* this call renders sample indices [sample_first, sample_first+sample_count) */
    int32_t  sample_count;   /* <= 0 means "all spp"                                                 */
    int32_t  max_depth;      /* MAX_DEPTH = 50 (main.cpp:32); <= 0 selects 50                        */
    float    t_min;          /* 0.001f (main.cpp:45); <= 0 selects 0.001f                            */
    uint64_t seed;           /* Philox key; the image depends only on (seed, pixel, sample index)    */
    int32_t  rng_mode;       /* 0 = reference-distribution (glm/rand() byte quirk, random.inl:19-27), 1 = plain uniform */
    int32_t  flags;          /* HRT_RENDER_* bits                                                    */
} hrt_render_params;

#define HRT_RENDER_MEGAKERNEL 1   /* one thread per path instead of the persistent wavefront (debug/validation) */
#define HRT_RENDER_WAVEFRONT_V1 2  /* every camera ray goes through the queues (the r1b/r1c pipeline): only in libraries built with
                                      -DHRT_WITH_V1_PIPELINE (A/B builds); the product ignores the flag */
#define HRT_RENDER_DEFOCUS 4       /* opt-in: the lens sampling camera.h:34 has commented out ("TODO: Add back in randomness"):
                                      rd = glm::circularRand(lensRadius), offset = u*rd.x + v*rd.y, lensRadius = aperture/2
                                      (camera.h:29).  Without the flag - and with aperture == 0 - the offset is zero like
                                      the reference's; the lens draw uses a spare word of the camera-jitter Philox call,
                                      so no other random number of the path changes. */

#define HRT_RENDER_RR 8            /* opt-in Russian roulette: the reference has none (main.cpp:43-76 walks every path to MAX_DEPTH or
                                      a miss).  From the third segment on a path survives its scatter with probability
                                      p = clamp(max(attenuation), 0.05, 1) and carries attenuation / p: same expectation, fewer rays per
                                      sample on closed scenes.  Own Philox slot: with the flag off the film is bit-identical to before. */

typedef struct hrt_stats {
    uint64_t rays;            /* closest-hit queries issued (= world->hit calls of main.cpp:45)      */
    uint64_t paths;           /* camera samples                                                      */
    uint64_t launches;        /* kernel launches of this library since the last reset                */
    double   ms_trace;        /* device time in the traversal kernel (CUDA events)                   */
    double   ms_shade;        /* device time in the shading / regeneration kernel                    */
    double   ms_other;        /* raygen, film, resets                                                */
    double   ms_total;        /* device time of the whole call                                       */
    uint64_t film_clamped;    /* finite sample channels beyond the film's +-2^21 range (clamped before the integer add;
                                 a reference float sum would not clamp): 0 on every BASELINE config (tested)              */
    uint64_t prim_tests;      /* rays answered by the last-resort reference-order walk (NaN t, zero direction components) */
    uint64_t retraced;        /* rays whose answer depended on the visiting order and was settled by replaying the
                                 reference's sequential process over the collected candidates (HRT_BVH_SAH only) */
    uint64_t scene_bytes;     /* bytes of scene data resident in device memory after hrt_scene_commit  */
    double   ms_build;        /* time of the BVH builds in hrt_scene_commit (host builders, or upload + device build)   */
    uint64_t h2d_bytes;       /* bytes hrt_scene_commit copied host -> device (scene arrays + mesh sources of device builds) */
    double   ms_build_upload; /* HRT_BVH_LBVH_GPU: the part of ms_build spent copying mesh sources to the device       */
} hrt_stats;

typedef struct hrt_scene hrt_scene;

/* ---- library ------------------------------------------------------------------------------- */
int         hrt_abi_version(void);
const char* hrt_last_error(void);                /* thread-local, never NULL                         */
int         hrt_device_count(void);              /* number of usable CUDA devices, < 0 on failure    */
int         hrt_set_device(int device);          /* device used by scenes created afterwards         */

/* ---- scene builder (one call per construction site of scene.cpp) ---------------------------- */
hrt_scene*  hrt_scene_create(void);
void        hrt_scene_destroy(hrt_scene* s);         /* parks the scene's device work buffers (ray queues, film) for the next
                                                        scene of this process instead of freeing them                      */
void        hrt_trim(void);                          /* frees what hrt_scene_destroy parked                                */

int hrt_scene_add_texture_solid(hrt_scene* s, const float rgb[3]);                    /* texture.h:9-25    */
int hrt_scene_add_texture_checker(hrt_scene* s, int even_tex, int odd_tex);          /* texture.cpp:17-28 */
int hrt_scene_add_texture_image(hrt_scene* s, const uint8_t* rgb8, int w, int h);    /* texture.cpp:53-74; w*h*3 bytes; w==0 -> cyan */
int hrt_scene_add_texture_env(hrt_scene* s, const float* data, int w, int h, int channels); /* texture.cpp:76-97 */
int hrt_scene_add_material(hrt_scene* s, const hrt_material* m);

/* shapes: return a shape id.  Primitive ids of a shape are hrt_scene_prim_base(shape) + local index. */
int hrt_scene_add_sphere(hrt_scene* s, const float center[3], float radius, int material);        /* sphere.h:9      */
int hrt_scene_add_rect(hrt_scene* s, int axis, float a0, float a1, float b0, float b1, float k, int material);
      /* axis 0: YZRect(y0,y1,z0,z1,k)  1: XZRect(x0,x1,z0,z1,k)  2: XYRect(x0,x1,y0,y1,k)      aarect.h          */
int hrt_scene_add_box(hrt_scene* s, const float center[3], const float dimensions[3], int material); /* box.h:41-44; 6 prims in box.h:31-38 order */
int hrt_scene_add_mesh(hrt_scene* s, const float* positions /* n_tri*9 */, const float* normals /* n_tri*9 or NULL */,
                       const float* uvs /* n_tri*6 or NULL */, int64_t n_tri, int material);       /* mesh.cpp:13-41: one ITriangle per face + BVH */
/* Mesh ingest fast path: the vertex / index buffers as Mesh::assimpLoadFile concatenates them over the file's aiMeshes
 * (mesh.cpp:63-115), flattened by the library instead of by the caller - Mesh::Mesh's loop (mesh.cpp:24-33): triangle k
 * takes vertices[indices[3k]], [3k+1], [3k+2] together with their normals and uvs (zeros when the file has none,
 * mesh.cpp:87-99).  With HRT_BVH_LBVH_GPU the compact buffers are what crosses PCIe and the gather runs on the device.
 * submeshes == NULL: indices are used as they are - the reference's behaviour, including its missing base-vertex offset
 * for files with several aiMeshes (mesh.cpp:111-114).  With a table (one entry per aiMesh) index j of sub-mesh m is
 * indices[first_index + j] + base_vertex: the corrected behaviour.  n_indices must be a multiple of 3 and every resolved
 * index < n_vertices (the reference reads out of bounds there; this library returns HRT_ERR_ARG).
 * The primitive ids, arrays and images are identical to hrt_scene_add_mesh on the flattened triangles.
 * Without HRT_MESH_BORROW the arrays are copied and may be freed as soon as the call returns.                             */
typedef struct hrt_submesh {
    int64_t first_index;   /* position of the sub-mesh's first index in `indices`             */
    int64_t n_index;       /* 3 * mNumFaces                                                    */
    int64_t base_vertex;   /* position of the sub-mesh's first vertex in `vertices`            */
} hrt_submesh;
int hrt_scene_add_mesh_indexed(hrt_scene* s, const float* vertices /* n_vertices*3 */, const float* normals /* n_vertices*3 or NULL */,
                               const float* uvs /* n_vertices*2 or NULL */, int64_t n_vertices,
                               const uint32_t* indices, int64_t n_indices,
                               const hrt_submesh* submeshes /* or NULL */, int n_submeshes, int material, uint32_t flags);
#define HRT_MESH_BORROW 1u   /* the caller keeps vertices / normals / uvs (and, without a sub-mesh table, indices) alive and unchanged
                                until hrt_scene_commit has returned; the library reads them in place instead of copying them at add
                                time (ownership rule of Scene's vectors: they outlive the load, mesh.cpp:15-21).  After the commit
                                the library no longer touches them. */
int64_t hrt_scene_prim_base(const hrt_scene* s, int shape);
int64_t hrt_scene_prim_count(const hrt_scene* s);

/* world list entries, in order (scene.cpp:356 objects.add).  ops are listed innermost first, i.e. in the
 * order the loader wraps them (rotate, scale, translate; scene.cpp:335-353). */
int hrt_scene_add_object(hrt_scene* s, int shape, const hrt_xform* ops, int n_ops);
/* ConstantMedium(boundary = wrapped shape, density, albedo)  constantMedium.h:10-14.  albedo_tex < 0 -> rgb.
 * The medium is one primitive (its id = hrt_scene_prim_base of the returned pseudo-shape id). */
int hrt_scene_add_medium(hrt_scene* s, int shape, const hrt_xform* ops, int n_ops, float density,
                         const float rgb[3], int albedo_tex);
int hrt_scene_set_background(hrt_scene* s, int texture);                                           /* scene.cpp:215-237 */

/* Upload + build acceleration structures.  seed drives the split-axis stream of HRT_BVH_REFERENCE. */
int hrt_scene_commit(hrt_scene* s, int bvh_mode, uint64_t seed);

/* ---- hot path ------------------------------------------------------------------------------ */
/* Closest-hit on a ray batch: world->hit(r, t_min, t_max, rec) of main.cpp:45 / hittableList.cpp:4-21.
 * rays: n x 8 floats (ox,oy,oz,t_min, dx,dy,dz,t_max).  Outputs: primitive id (HRT_MISS on miss) and rec.t.
 * optional outputs (may be NULL): normal n x 3 (rec.normal), uv n x 2, front n bytes (rec.frontFace).
 * rng_seed only matters for scenes with media (their hit() draws a random number).
 * Batches of 4 Mi rays and more without the optional outputs (scenes without media) run in 2 Mi-ray chunks on internal streams, the copies of
 * one chunk overlapping the walks of another: give pinned host memory to let them overlap; the answers do not depend on the chunking. */
int hrt_trace_batch(hrt_scene* s, const float* rays, int64_t n, uint32_t* prim_id, float* t,
                    float* normal, float* uv, uint8_t* front, uint64_t rng_seed);
/* Same with device-resident buffers (inputs already in HBM); stream = cudaStream_t or NULL. */
int hrt_trace_batch_device(hrt_scene* s, const void* d_rays, int64_t n, void* d_prim_id, void* d_t,
                           uint64_t rng_seed, void* stream);

/* hrt_trace_batch_device for callers that must not allocate device memory (CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  No counterpart in the reference.  Same inputs, outputs
 * and results as hrt_trace_batch_device, but the scratch of the batch kernels (16 B hit record + 4 B pending-list slot per ray, 256 B
 * of control words) is a file-scope device array of HRT_TRACE_BATCH_STATIC_MAX rays (20 MiB + 256 B, held by every GPU that loads the
 * library until the process exits) instead of cudaMalloc memory.  The call never calls cudaMalloc / cudaFree or an allocating library
 * routine; its first use on a device looks the array's address up (cudaGetSymbolAddress) - make that call outside a graph capture.
 * n > HRT_TRACE_BATCH_STATIC_MAX is HRT_ERR_ARG (split the batch).  All calls - of all scenes on that GPU - share the array: the
 * control words are cleared by every call, and calls that overlap on the GPU (different streams, no ordering) race.               */
#define HRT_TRACE_BATCH_STATIC_MAX (1 << 20)
int hrt_trace_batch_static(hrt_scene* s, const void* d_rays, int64_t n, void* d_prim_id, void* d_t,
                           uint64_t rng_seed, void* stream);

/* The shading half of the path on caller-supplied batches (parity entries, like hrt_trace_batch for hit()):
 * hrt_texture_eval: Texture::colourValue(u, v, p) of texture `texture` (texture.h:3-7, texture.cpp:17-97); uvp: n x 5 floats
 *   (u, v, p.xyz), out_rgb: n x 3.
 * hrt_scatter_batch: Material::emitted + Material::scatter of material `material` (material.h:60-242, material.cpp:18-28) on n hit
 *   records; in14: n x 14 floats (r_in.origin, r_in.dir, rec.p, rec.normal, rec.u, rec.v), front: n bytes (rec.frontFace); out10:
 *   n x 10 floats (scattered ? 1 : 0, attenuation rgb, scattered.dir, emitted rgb).  Record i draws from the Philox stream of
 *   (seed; index i, sample 0) at bounce 0, rng_mode as in hrt_render_params. */
/* hrt_camera_rays: Camera::getRay(s, t) of Camera(cam) (camera.h:9-39; the lens offset is zero like the reference's) on n (s, t) pairs, computed on
 * the device: out = n x 6 floats (origin xyz, direction xyz - not normalised).  Needs no scene; runs on the device of hrt_set_device. */
int hrt_camera_rays(const hrt_camera* cam, const float* st, int64_t n, float* out);
int hrt_texture_eval(hrt_scene* s, int texture, const float* uvp, int64_t n, float* out_rgb);
int hrt_scatter_batch(hrt_scene* s, int material, const float* in14, const uint8_t* front, int64_t n, uint64_t seed, int rng_mode, float* out10);

/* render(): main.cpp:81-140.  out_sum: width*height*3 floats receiving the per-pixel *sum* of the rendered
 * samples (NULL to skip); out_rgb8: width*height*3 bytes = Film::tonemap + writeColour of sum/spp
 * (film.cpp:25-52), row 0 = top row as in the reference's pixel buffer (NULL to skip).
 * Device memory besides the scene: the integer film (32 B per pixel) and the ray queues of the wavefront, 148 B per entry for
 * min(32 Mi, camera paths of the call) entries (at most ~5 GB; HRT_WAVE_CAPACITY overrides the entry count); both are kept for the next
 * render and for the next scene of the process (hrt_trim releases them). */
int hrt_render(hrt_scene* s, const hrt_camera* cam, const hrt_render_params* p, float* out_sum, uint8_t* out_rgb8);
/* render() in slices, with checkpoints and the reference's progress reporting (main.cpp:95-109 prints "Pixels rendered"
 * from a reporter thread; here the caller gets a callback after every slice).  The p->spp samples of every pixel are
 * rendered slice_spp at a time (<= 0: one slice).  checkpoint_path != NULL: after every slice the integer film is written
 * there (temporary file + rename) together with the render parameters; when the file already exists and belongs to the same
 * render (size, spp, seed, rng mode, depth, t_min, camera, primitive count) rendering RESUMES after the samples it holds,
 * and a file from a different render is HRT_ERR_IO, never silently overwritten.  The film is integer and Philox is keyed
 * by the sample index, so the image is bit-identical to hrt_render's for any slicing and any number of interruptions.
 * on_slice (may be NULL) is called on the calling thread with the samples finished so far; returning non-zero stops the
 * render: the call returns HRT_STOPPED (1) with the checkpoint kept and no image written.  p->sample_first/sample_count
 * must select the whole image (0 / <= 0 or spp).                                                                        */
#define HRT_STOPPED 1
typedef int (*hrt_progress_fn)(int samples_done, int spp, void* user);
int hrt_render_progressive(hrt_scene* s, const hrt_camera* cam, const hrt_render_params* p, const char* checkpoint_path,
                           int slice_spp, hrt_progress_fn on_slice, void* user, float* out_sum, uint8_t* out_rgb8);
/* The checkpoint file on its own (tooling): samples held by a checkpoint, or HRT_ERR_IO.                                 */
int hrt_checkpoint_samples(const char* checkpoint_path, int* width, int* height, int* spp);

/* render() on several GPUs of one process: main.cpp:176 stays ONE synchronous call.  `devices` lists n_devices distinct CUDA
 * device ordinals, devices[0] being the one the scene was committed on.  The committed scene is replicated once per extra device
 * (uploads of the same host arrays; device-built meshes are copied over NVLink) and kept for later calls; one host thread per
 * device renders a contiguous share of the p->spp sample indices into that device's integer film, and devices[0] sums the films
 * in place over peer memory and resolves them in one kernel (no collective, no extra process).  The image is bit-identical to
 * hrt_render's for any number of devices (integer film, Philox keyed by the global sample index).  p->sample_first / sample_count
 * must select the whole image. */
int hrt_render_multi(hrt_scene* s, const int* devices, int n_devices, const hrt_camera* cam, const hrt_render_params* p,
                     float* out_sum, uint8_t* out_rgb8);

/* Device variants for multi-GPU use.  The FILM is a device buffer of hrt_film_bytes(width, height) bytes:
 * per pixel four int64 lanes — the r, g, b sample sums in fixed point (2^-26 units; every sample rounded to
 * nearest and clamped to +-2^21) and one lane of "a NaN / +inf / -inf sample was added" fields.  Integer sums
 * commute: the film does not depend on the order in which paths finish nor on how the samples were split over
 * GPUs, so an integer SUM reduction of the films of all ranks (ncclSum on int64) gives bit-identical images for
 * any GPU count.  The caller zeroes the film; hrt_render_accumulate_device adds sample indices
 * [sample_first, sample_first + sample_count) of every pixel; hrt_film_resolve_device turns a (reduced) film
 * into float sums / float means (sum / spp, main.cpp:126) / the 8-bit image (film.cpp:25-52), any of which
 * may be NULL.  `stream` is a cudaStream_t; NULL selects the scene's own (non-blocking) stream, and the call then first waits for
 * whatever the caller queued on the legacy default stream (hrt_film_zero with NULL is complete when it returns). */
int64_t hrt_film_bytes(int width, int height);
int hrt_render_accumulate_device(hrt_scene* s, const hrt_camera* cam, const hrt_render_params* p,
                                 void* d_film, void* stream);
int hrt_film_resolve_device(hrt_scene* s, const void* d_film, int width, int height, int spp,
                            void* d_sum_or_null, void* d_mean_or_null, void* d_rgb8_or_null, void* stream);

/* Reduce + resolve over NVLink peer memory (one process per GPU).  Films that other processes can map are allocated by the
 * library (hrt_film_create: a zeroed cudaMalloc allocation of hrt_film_bytes, addressable by a CUDA IPC handle).  Every rank
 * renders into its own film and publishes hrt_film_ipc_handle; after a barrier the root calls hrt_film_resolve_peers with the
 * handles of the OTHER ranks: one kernel pulls the peers' lanes over NVLink, adds them to the local film as integers (the
 * same bits as an ncclSum reduce) and writes float sums / means / the 8-bit image; write_back != 0 also stores the reduced
 * film.  The peers must not touch their films until the root's stream has finished (barrier after the call). */
#define HRT_IPC_HANDLE_BYTES 64
#define HRT_MAX_PEERS 15
void* hrt_film_create(int width, int height);
void  hrt_film_destroy(void* d_film);
int   hrt_film_zero(void* d_film, int width, int height, void* stream);
int   hrt_film_ipc_handle(void* d_film, void* out_handle);
int   hrt_film_resolve_peers(hrt_scene* s, void* d_film, const void* peer_handles, int n_peers, int width, int height, int spp,
                             int write_back, void* d_sum_or_null, void* d_mean_or_null, void* d_rgb8_or_null, void* stream);

/* The same reduction spread over the ranks (what bench.py --gpus N and hobbyraytracer_b200/multigpu.py use).  Buffers that
 * other processes map are library allocations: hrt_peer_alloc (zeroed cudaMalloc memory on the current device), hrt_peer_handle (its
 * CUDA IPC handle, HRT_IPC_HANDLE_BYTES), hrt_peer_open (maps another process's handle into this one; cached per scene and closed
 * with it).  Every rank holds its own film and the mappings of all other films and of the ROOT's output buffers; after the renders
 * (a stream-ordered barrier between them and the reduction is the caller's: e.g. a one-word ncclAllReduce on the same stream) every
 * rank calls hrt_film_resolve_slice on ITS share of the pixels: the lanes of all `films` (own first or not: integer sums commute) are
 * added and resolved, and float sums / means / the 8-bit image of pixels [pix_first, pix_first + pix_count) are stored through the
 * given pointers - the root's buffers - at their pixel offsets.  No rank gathers alone, nothing is staged; a second barrier tells the
 * root that every slice has landed and the ranks that their films may be zeroed again. */
void* hrt_peer_alloc(int64_t bytes);
void  hrt_peer_free(void* p);
int   hrt_peer_handle(void* p, void* out_handle);
void* hrt_peer_open(hrt_scene* s, const void* handle);
int   hrt_film_resolve_slice(hrt_scene* s, const void* const* films, int n_films, int width, int height, int spp, int pix_first, int pix_count,
                             void* d_sum_or_null, void* d_mean_or_null, void* d_rgb8_or_null, void* stream);

/* Introspection of a committed scene (tests, tooling): copies one device array of mesh `mesh_index` (world-list order of
 * first use) into `out` (may be NULL) and returns its size in bytes.  which: 0 reference-topology nodes (32 B each),
 * 1 fast-tree nodes (64 B), 2 leaf records (128 B), 3 triangle positions (48 B), 4 triangle attributes (64 B),
 * 5 nodes of the 8-wide tree the render kernels walk (96 B). */
int64_t hrt_scene_mesh_array(hrt_scene* s, int mesh_index, int which, void* out, int64_t capacity_bytes);

int hrt_get_stats(hrt_scene* s, hrt_stats* out);
int hrt_reset_stats(hrt_scene* s);

#ifdef __cplusplus
}
#endif
#endif /* HRT_H */
