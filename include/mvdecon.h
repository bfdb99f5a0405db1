/*
 * mvdecon.h -- C ABI of libmvdecon.so, the B200 (sm_100a) implementation of the multi-view
 * Richardson-Lucy deconvolution inner loop of PreibischLab/SPIM_Registration.
 *
 * Part 1 are the symbols the reference binds through JNA today; a library exporting them is a
 * drop-in for the reference's "FourierConvolutionCUDALib".  Part 2 is the resident whole-loop API
 * (SURVEY.md section 8b "Level-1"), which keeps every view, weight, kernel spectrum, psi and the
 * quotient in HBM for the whole deconvolution.  Citations are relative to the reference tree,
 * J/ = src/main/java/.
 *
 * Conventions: all pixel data float32, x fastest (index = x + X*(y + Y*z)).  Part 1 passes
 * dimensions reversed (z,y,x) exactly as the reference does (MVDeconFFTThreads.java:136-144);
 * Part 2 passes dimensions as (x,y,z).  No function retains a caller pointer after it returns.
 * No function calls exit()/abort(); Part 2 functions return 0 on success, non-zero on failure,
 * and every failure latches a message readable with mvdGetLastError().
 */
#ifndef MVDECON_H
#define MVDECON_H

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------------------------------
 * Part 1 -- legacy JNA symbols
 * ------------------------------------------------------------------------------------------ */

/* J/spim/process/cuda/CUDAStandardFunctions.java:15 */
int getCUDAcomputeCapabilityMinorVersion(int devCUDA);
/* CUDAStandardFunctions.java:16 */
int getCUDAcomputeCapabilityMajorVersion(int devCUDA);
/* CUDAStandardFunctions.java:17-20 : -1 = driver failure, 0 = no device (CUDATools.java:37-48) */
int getNumDevicesCUDA(void);
/* CUDAStandardFunctions.java:21 : caller passes a zeroed 256-byte buffer (CUDATools.java:55,63-69) */
void getNameDeviceCUDA(int devCUDA, char* name);
/* CUDAStandardFunctions.java:22 : total device memory in bytes (Java long) */
long long getMemDeviceCUDA(int devCUDA);
/* CUDAStandardFunctions.java:23 : free device memory in bytes (optional in the reference,
 * CUDATools.java:76-84 catches UnsatisfiedLinkError) */
long long getFreeMemDeviceCUDA(int devCUDA);

/* J/spim/process/cuda/CUDAFourierConvolution.java:10, called from
 * MVDeconFFTThreads.java:65-68 and :88-91 (and legacy LRFFTThreads.java:49,64).
 * Convolves the block `im` (imDim = {z,y,x}) in place with `kernel` (kernelDim = {z,y,x}), kernel
 * centre at floor(K/2) (J/fftMethods/FFTConvolution.java:506-516).  Values outside the block are
 * taken as zero; the reference only consumes block coordinates [K/2, B-1-K/2]
 * (Block.java:313-359), for which the extension is irrelevant.  Re-entrant: one host thread per
 * device may call concurrently (MVDeconFFT.java:426-445).  On failure `im` is left untouched and
 * the error is latched (mvdGetLastError). */
void convolution3DfftCUDAInPlace(float* im, int* imDim, float* kernel, int* kernelDim, int devCUDA);

/* spim.process.cuda.CUDASeparableConvolution (CUDASeparableConvolution.java:13-17), the reference's
 * second native hook (Gaussian blurs of the Difference-of-Gaussian detection,
 * CUDASeparableConvolutionFunctions.java:127-199): in-place separable convolution with one kernel of
 * diameter N per axis, image x fastest.  outofbounds 0 = zero, 1 = outofboundsvalue, 2 = extend the
 * last pixel.  Java `boolean` arguments and results are C ints (JNA).  Returns 1 on success, 0 on
 * failure (mvdGetLastError).  N <= 31 runs the direct separable kernels, 63 and 127 the FFT engine axis by
 * axis (127 taps along x or y use the 256-point windows; called at CUDASeparableConvolutionFunctions.java:191). */
int convolve_7(float* image, float* kernelX, float* kernelY, float* kernelZ, int imageW, int imageH, int imageD, int convolveX,
               int convolveY, int convolveZ, int outofbounds, float outofboundsvalue, int devCUDA);
int convolve_15(float* image, float* kernelX, float* kernelY, float* kernelZ, int imageW, int imageH, int imageD, int convolveX,
                int convolveY, int convolveZ, int outofbounds, float outofboundsvalue, int devCUDA);
int convolve_31(float* image, float* kernelX, float* kernelY, float* kernelZ, int imageW, int imageH, int imageD, int convolveX,
                int convolveY, int convolveZ, int outofbounds, float outofboundsvalue, int devCUDA);
int convolve_63(float* image, float* kernelX, float* kernelY, float* kernelZ, int imageW, int imageH, int imageD, int convolveX,
                int convolveY, int convolveZ, int outofbounds, float outofboundsvalue, int devCUDA);
int convolve_127(float* image, float* kernelX, float* kernelY, float* kernelZ, int imageW, int imageH, int imageD, int convolveX,
                 int convolveY, int convolveZ, int outofbounds, float outofboundsvalue, int devCUDA);
/* CUDASeparableConvolution.convolutionCPU (CUDASeparableConvolution.java:19-21; declared by the reference's interface, called
 * nowhere in it): in-place separable convolution with kernels of RADIUS kernelR* (2 R + 1 taps, NULL = leave the axis alone),
 * outofbounds as above.  This library has no CPU code: the result is computed on device 0 by the kernels behind convolve_N. */
void convolutionCPU(float* image, float* kernelX, float* kernelY, float* kernelZ, int kernelRX, int kernelRY, int kernelRZ, int imageW,
                    int imageH, int imageD, int outofbounds, float outofboundsvalue);

/* CUDAFourierConvolution.java:9 -- declared by the reference, never called.  Returns a
 * malloc()ed array with the result (caller frees) or NULL on failure. */
float* convolution3DfftCUDA(float* im, int* imDim, float* kernel, int* kernelDim, int devCUDA);

/* ------------------------------------------------------------------------------------------
 * Part 2 -- resident deconvolution loop
 * Mirrors MVDeconvolution (J/spim/process/fusion/deconvolution/MVDeconvolution.java:73-190,
 * 333-444) with MVDeconFFT.convolve1/convolve2 (MVDeconFFT.java:363-535) fused in.
 * ------------------------------------------------------------------------------------------ */

typedef struct mvdContext mvdContext;

/* Copies the last error message of the calling thread (or of the last failing legacy call) into
 * buf (n bytes, always NUL-terminated).  Returns the latched code (0 = no error) and clears it. */
int mvdGetLastError(char* buf, int n);

/* Library/build description ("mvdecon-b200 <version> sm_100a ...") */
const char* mvdVersion(void);

#define MVD_MEM_HOST 0
#define MVD_MEM_DEVICE 1
#define MVD_MEM_HOST_IMG_DEVICE_WEIGHT 2 /* mvdAddView only: img on the host, weight evaluated on the device (mvdVirtualWeight) */
#define MVD_MEM_HOST_ASYNC 3 /* mvdAddView only: host arrays (pinned for a real overlap) that the caller keeps alive until the next
                              * mvdSynchronize / mvdRunIterations / mvdInitPsiAverage returns: the call does not wait for the upload, the
                              * first step that reads the view does (on the device) */
#define MVD_MEM_HOST_ASYNC_IMG_DEVICE_WEIGHT 4 /* mvdAddView only: img as MVD_MEM_HOST_ASYNC, weight on the device (copied before return) */

/* Create a context on CUDA device `device` for one brick of the deconvolved volume.
 *   globalDim   (x,y,z) size of the whole deconvolved volume psi.
 *   brickOffset (x,y,z) first voxel of this brick, brickDim its size; NULL, NULL = whole volume.
 *   ghost       (x,y,z) halo width kept around the brick for psi and the quotient (filled by the
 *               caller's halo exchange, see mvdArrayInfo); NULL or zeros for a single brick.  Ghost
 *               layers are clipped at the global volume border, where the reference's
 *               out-of-bounds rules apply instead (mirror-single for psi, constant 1 for the
 *               quotient; MVDeconFFTThreads.java:23,42,57,82).
 *   lambda      Tikhonov parameter, 0 = off (EfficientBayesianBased.java:89,252-253).
 *   minValue    floor of psi, 1e-4f in the reference (MVDeconvolution.java:49). */
int mvdCreate(mvdContext** ctx, int device, const int globalDim[3], const int brickOffset[3], const int brickDim[3],
              const int ghost[3], double lambda, float minValue);
int mvdDestroy(mvdContext* ctx);

/* Use an existing CUDA stream (cudaStream_t passed as void*) for all work of this context. */
int mvdSetStream(mvdContext* ctx, void* cudaStream);
int mvdSynchronize(mvdContext* ctx);

/* Add one view (MVDeconInput.add / new MVDeconFFT, MVDeconFFT.java:57-158).
 *   img      brickDim floats: the transformed view, 0 = no data (MVDeconvolution.java:496).
 *   weight   brickDim floats or NULL; with NULL the constant `weightConst` is used (the
 *            reference's NO_WEIGHTS case after normalisation, ProcessForDeconvolution.java:351-384).
 *   kernel1, kernel2  already prepared as by MVDeconFFT.init (MVDeconFFT.java:162-303), dims (x,y,z).
 *   memKind  MVD_MEM_HOST or MVD_MEM_DEVICE: where img and weight live (kernels are always host);
 *            MVD_MEM_HOST_IMG_DEVICE_WEIGHT: img on the host, weight on the device.
 * Returns the view index through *viewIndex (may be NULL). */
int mvdAddView(mvdContext* ctx, const float* img, const float* weight, float weightConst, const float* kernel1,
               const int kernel1Dim[3], const float* kernel2, const int kernel2Dim[3], int memKind, int* viewIndex);
int mvdNumViews(mvdContext* ctx);

/* psi initialisation (MVDeconvolution.java:105-127, FirstIteration.java:63-171):
 * mvdFirstIterationSums returns this brick's sum of per-voxel view averages and the count of
 * voxels with data; mvdFillPsi sets psi (brick and ghost) to a constant; mvdInitPsiAverage does
 * both for a single brick (NaN average -> 0.5) and returns the average. */
int mvdFirstIterationSums(mvdContext* ctx, double* sum, long long* count);
int mvdFillPsi(mvdContext* ctx, float value);
int mvdInitPsiAverage(mvdContext* ctx, double* avg);
/* Start from a given psi instead (MVDeconvolution.initialImage, MVDeconvolution.java:95-104):
 * brickDim floats; values <= 0 or NaN are not touched here (the reference replaces them while
 * loading, MVDeconvolution.java:237-320 -- do that before the call). */
int mvdSetPsi(mvdContext* ctx, const float* psi, int memKind);
int mvdGetPsi(mvdContext* ctx, float* psi, int memKind);

/* Whole iterations on a single brick (MVDeconvolution.runIteration, :333-444): for every view in
 * order conv1 -> quotient -> conv2 -> update.  sumChange/maxChange (may be NULL) receive
 * numIterations*numViews values, index = iteration*numViews + view (the per-view log line of
 * MVDeconvolution.java:427-436).  Fails if the context has ghost layers (use the step API). */
int mvdRunIterations(mvdContext* ctx, int numIterations, double* sumChange, double* maxChange);

/* Step API for bricks with ghost layers.  Per view and iteration the caller runs
 *   mvdStepQuotient(v)   quotient = img_v > 0 ? img_v / conv1_v(psi) : 1        on the brick
 *   <exchange MVD_ARRAY_QUOTIENT ghost layers with the neighbouring bricks>
 *   mvdStepUpdate(v)     psi = update(psi, conv2_v(quotient), w_v, lambda)       on the brick
 *   <exchange MVD_ARRAY_PSI ghost layers>
 * statSlot indexes a device-side statistics table read back with mvdGetStats. */
int mvdStepQuotient(mvdContext* ctx, int view);
int mvdStepUpdate(mvdContext* ctx, int view, int statSlot);
int mvdGetStats(mvdContext* ctx, int firstSlot, int numSlots, double* sumChange, double* maxChange);

/* Per-kernel timing with CUDA events on the context's stream, for bench.py's roofline: between
 * Start and Stop every launch of the three convolution kernels is bracketed by an event pair.
 * Stop returns, per kernel (0 fwd_xy, 1 z_pass, 2 inv_xy), the summed duration in ms, the launch
 * count and the summed ALGORITHMIC bytes (DESIGN.md byte model). */
int mvdProfileStart(mvdContext* ctx);
int mvdProfileStop(mvdContext* ctx, double ms[3], long long launches[3], double bytes[3]);

#define MVD_ARRAY_PSI 0
#define MVD_ARRAY_QUOTIENT 1
/* Device pointer and geometry of a ghosted array: dim = extent, off = index of brick voxel 0. */
int mvdArrayInfo(mvdContext* ctx, int which, void** devicePtr, int dim[3], int off[3]);

/* Direct path for small separable kernels.  mvdAddView tests every kernel for rank-1 separability
 * (sum |K - a(z) b(y) c(x)| <= 1e-6 sum |K|, each extent <= 31); such convolutions run as direct
 * separable passes with the same fused epilogues instead of through the FFT engine (Gaussian PSFs
 * and the compound kernels built from them qualify, extracted bead PSFs do not).  Enabled by
 * default, also for bricks with attached peers (the z kernel's epilogue pushes ghost voxels like the FFT path's).  The
 * substitution is an approximation the reference does not make: the factors are the kernel's marginal sums, the
 * replaced kernel differs from the given one by at most 1e-6 of its L1 norm, i.e. the convolution result by at most
 * 1e-6 relative -- two orders of magnitude inside the 1e-4 parity bound; mvdSetDirectPath(ctx, 0) switches it off
 * (every convolution then runs the FFT engine on the kernel as given).  mvdViewUsesDirectPath reports the tap
 * count class (0 = FFT, else 7 / 15 / 31) chosen for conv1 and conv2 of a view. */
int mvdSetDirectPath(mvdContext* ctx, int enable);
int mvdViewUsesDirectPath(mvdContext* ctx, int view, int* conv1, int* conv2);

/* Fused ghost exchange.  Instead of exchanging ghost layers between the two steps, a context can
 * be told where its neighbours' ghosted arrays live; mvdStepQuotient / mvdStepUpdate then store
 * every output voxel a neighbour needs straight into that neighbour's ghost layers from the
 * kernel that computes it (peer memory over NVLink), and the "exchange" shrinks to a barrier
 * between the steps (all bricks finished step n before any starts step n+1).
 *   mvdExportArrays  64-byte CUDA IPC handles of this context's psi and quotient arrays, to be
 *                    sent to the processes that own the neighbouring bricks.
 *   mvdAttachPeer    one neighbour (any brick whose ghost layers overlap this brick): its
 *                    brickOffset / brickDim as given to its mvdCreate (same globalDim and ghost),
 *                    and its arrays as MVD_PEER_POINTER (device pointers usable from this
 *                    context's device: same process, same device or peer access enabled by the
 *                    library) or MVD_PEER_IPC (the 64-byte handles from mvdExportArrays of another
 *                    process).  peerDevice is the CUDA device of the neighbour (MVD_PEER_POINTER).
 *   mvdDetachPeers   back to caller-side exchanges; closes IPC mappings.
 * The reference has no counterpart (one Java thread per device pulling blocks through host RAM,
 * MVDeconFFTThreads.java:96-134). */
#define MVD_PEER_POINTER 0
#define MVD_PEER_IPC 1
int mvdExportArrays(mvdContext* ctx, unsigned char psiHandle[64], unsigned char quotientHandle[64]);
int mvdAttachPeer(mvdContext* ctx, const int peerBrickOffset[3], const int peerBrickDim[3], const void* psi,
                  const void* quotient, int kind, int peerDevice);
int mvdDetachPeers(mvdContext* ctx);

/* Inputs of the loop produced on the device (the step before the loop).
 * mvdTransformView = TransformInputAndWeights (TransformInputAndWeights.java:56-114) for one view:
 *   src / srcDim      the view's image stack (x,y,z)
 *   inverse           row-major 3x4 matrix world -> view coordinates, i.e. the inverse of the view's
 *                     registration (AffineTransform3D.applyInverse, :100); world = output voxel + outOffset
 *   outOffset, outDim bounding box of the deconvolved volume (ProcessForDeconvolution.java:198)
 *   blendBorder/Range Blending parameters (ProcessForDeconvolution.java:545-559); used when outWeight != NULL
 *   outImg            max(minValue, n-linear interpolation over the mirror-single extended view) inside
 *                     the view, 0 outside (0 = no data); outWeight the cosine blending weight
 *                     (BlendingRealRandomAccess.java:71-101), may be NULL (TransformInput.java:50-95).
 * mvdNormalizeWeights = WeightNormalizer with directly applied normalisation (WeightNormalizer.java:
 * 133-181: every weight divided by the sum over the views, 0/0 stays NaN as in the reference) followed by
 * adjustForOSEM (ProcessForDeconvolution.java:351-362: w = min(1, w * osemSpeedup), skipped for 1.0);
 * returns the minimal and the average number of views with weight > 0 per voxel (:97-101). */
int mvdTransformView(int device, const float* src, const int srcDim[3], const double inverse[12], const int outOffset[3],
                     const int outDim[3], const float blendBorder[3], const float blendRange[3], float minValue, float* outImg,
                     float* outWeight, int srcMemKind, int outMemKind);
int mvdNormalizeWeights(int device, float* const* weights, int numViews, long long numVoxels, double osemSpeedup,
                        int* minOverlappingViews, double* avgOverlappingViews, int memKind);

/* Virtual weights -- the dialog's default weight type (EfficientBayesianBased.java:71,84;
 * ProcessForDeconvolution.java:204-208, 308-309, 375-379).  The reference keeps no weight images: every
 * read of w_v(x) inside the loop evaluates
 *     min(1, blend_v(A_v^-1 (x + outOffset)) / sumW(x) * osemSpeedup)        NormalizingRandomAccess.java:36-45
 * with the position from TransformedInterpolatedRealRandomAccess.applyInverse (:68-95) and the float image
 *     sumW(x) = sum_u blend_u(...) where that exceeds 1, else 1              WeightNormalizer.java:183-232
 * With 180 GB of HBM the same values are evaluated once and stay resident (pass them to mvdAddView as dense
 * weights): identical results, no affine transform and lookup per voxel, view and iteration.
 *   mvdVirtualWeightSum  sumW and the overlap statistics (WeightNormalizer.java:74-88) for numViews views:
 *                        srcDims 3 ints, inverses 12 doubles (row-major 3x4, world -> view), blendBorders /
 *                        blendRanges 3 floats per view.  avgOverlappingViews is the mean over all voxels (the
 *                        reference averages the means of 2 x numThreads equal portions).
 *   mvdVirtualWeight     one view's weight image from sumW and the OSEM speed-up
 *                        (NormalizingRandomAccessibleInterval.setOSEMspeedup, ProcessForDeconvolution.java:375-379).
 * sumWeights / outWeight: host or device arrays of outDim voxels (x fastest) according to memKind. */
int mvdVirtualWeightSum(int device, int numViews, const int* srcDims, const double* inverses, const float* blendBorders,
                        const float* blendRanges, const int outOffset[3], const int outDim[3], float* sumWeights,
                        int* minOverlappingViews, double* avgOverlappingViews, int memKind);
int mvdVirtualWeight(int device, const int srcDim[3], const double inverse[12], const float blendBorder[3], const float blendRange[3],
                     const int outOffset[3], const int outDim[3], const float* sumWeights, double osemSpeedup, float* outWeight,
                     int memKind);

/* PSF extraction (ExtractPSF.java).
 * mvdExtractPSF = extractPSFLocal (:353-394): psf(v) = sum over the bead locations (x,y,z triples, view
 *   coordinates) of the view sampled n-linearly with periodic extension at v - psfSize/2 + location, float
 *   accumulation in the order given; normalize != 0 applies ExtractPSF.normalize (:277-295, min-max to [0,1]).
 * mvdTransformedPSFSize / mvdTransformPSF = transformPSF (:304-346) + transform (:396-437): the PSF resampled
 *   (zero outside) through the view's model (row-major 3x4, view -> world) into an odd-sized image whose
 *   centre is the transformed centre voxel; newDim comes from mvdTransformedPSFSize (host only).
 * psf / out are host arrays (PSFs are small). */
int mvdExtractPSF(int device, const float* img, const int imgDim[3], const double* locations, int numLocations,
                  const int psfSize[3], int normalize, float* psf, int imgMemKind);
int mvdTransformedPSFSize(const int psfDim[3], const double model[12], int newDim[3]);
int mvdTransformPSF(int device, const float* psf, const int psfDim[3], const double model[12], const int newDim[3], float* out);

/* psi = 0 where no view has data (MVDeconvolution.java:180-187). */
int mvdMaskNeverUpdated(mvdContext* ctx);

/* ---- several devices behind one handle, one process -------------------------------------------------------------
 * The reference drives N devices from one JVM, one Java thread per device pulling blocks through host RAM
 * (MVDeconFFT.java:426-445, MVDeconFFTThreads.java:96-134).  mvdMulti keeps the same calling pattern -- one handle,
 * one process, no torch / NCCL -- but shards the resident state: the volume is cut into one brick per device
 * (power-of-two device count; longest axis first), every brick with ghost layers of the kernel half-width, and the
 * step kernels store ghost voxels straight into the neighbouring devices' arrays (peer access over NVLink).  A
 * worker thread per device issues its launches; between two steps a device waits only for the events of its
 * neighbours.  `devices` may name a device several times (bricks then share it).
 *   maxKernelDim  largest extent per axis of any kernel1 / kernel2 that will be added (sizes the ghost layers)
 *   img / weight of mvdMultiAddView and psi of mvdMultiSetPsi / mvdMultiGetPsi are WHOLE-VOLUME host arrays
 *   (x fastest, extent globalDim); statistics are summed / maximised over the bricks. */
typedef struct mvdMulti mvdMulti;
int mvdCreateMulti(mvdMulti** multi, const int* devices, int numDevices, const int globalDim[3], const int maxKernelDim[3],
                   double lambda, float minValue);
int mvdDestroyMulti(mvdMulti* multi);
int mvdMultiNumBricks(mvdMulti* multi);
int mvdMultiBrickInfo(mvdMulti* multi, int brick, int* device, int offset[3], int dim[3]);
int mvdMultiAddView(mvdMulti* multi, const float* img, const float* weight, float weightConst, const float* kernel1,
                    const int kernel1Dim[3], const float* kernel2, const int kernel2Dim[3]);
int mvdMultiSetDirectPath(mvdMulti* multi, int enable);
int mvdMultiInitPsiAverage(mvdMulti* multi, double* avg);
int mvdMultiSetPsi(mvdMulti* multi, const float* psi);
int mvdMultiRunIterations(mvdMulti* multi, int numIterations, double* sumChange, double* maxChange);
int mvdMultiMaskNeverUpdated(mvdMulti* multi);
int mvdMultiGetPsi(mvdMulti* multi, float* psi);

/* Kernel preparation behind the C ABI: MVDeconInput.init / MVDeconFFT.init
 * (MVDeconInput.java:43-49, MVDeconFFT.java:162-333) with AdjustInput.normImg (AdjustInput.java:29-100, its sumImg
 * double-count included), Mirror.mirror (Mirror.java:31-108, even-size quirk included) and the compound kernels of
 * the four PSFTYPEs; the small zero-extended convolutions run through the device engine.  The Java host keeps its own
 * MVDeconFFT.init; this entry is for hosts without it.
 *   psfs[v]      host array, extent psfDims[3 v .. 3 v + 2] = (x, y, z), x fastest; views are initialised in order
 *   psfType      0 OPTIMIZATION_II, 1 OPTIMIZATION_I, 2 EFFICIENT_BAYESIAN, 3 INDEPENDENT (order of MVDeconFFT.PSFTYPE, :28)
 *   numThreads   the thread count that decides sumImg's portions (the reference: Threads.numThreads())
 *   kernel1[v], kernel2[v]  host outputs of the extent of psfs[v] */
int mvdPrepareKernels(int device, int numViews, const float* const* psfs, const int* psfDims, int psfType, int numThreads,
                      float* const* kernel1, float* const* kernel2);

/* Neighbour-only barrier of the fused exchange, in peer memory (one process per GPU: events cannot be shared, and a
 * collective over all ranks costs ~25 us per step).  Every context owns a small flag array; after a step a rank stores the
 * step number into its slot of every neighbour's array, before the next step it waits until all its neighbours have done
 * the same (ld.acquire.sys spin, bounded to ~10 s: on expiry the next mvdSynchronize / mvdGetStats fails).
 *   mvdExportFlags      the CUDA IPC handle and / or the device pointer of this context's flag array
 *   mvdAttachPeerFlags  one neighbour's array (MVD_PEER_IPC handle or MVD_PEER_POINTER), the slot (0 .. 31) of THIS brick in
 *                       the neighbour's array and the slot of the neighbour in this brick's array (e.g. the 3 x 3 x 3 offset code)
 *   mvdPeerSignal / mvdPeerWait / mvdPeerBarrier(ctx, value)  enqueue on the context's stream; values must increase by at most
 *                       2^31 per call.  mvdPeerBarrier = signal + wait in one kernel: only when every brick has its OWN GPU --
 *                       kernels of several bricks on one GPU that wait for one another may never run at the same time; there,
 *                       signal for all bricks first and wait afterwards. */
int mvdExportFlags(mvdContext* ctx, unsigned char flagsHandle[64], void** flagsPointer);
int mvdAttachPeerFlags(mvdContext* ctx, const void* peerFlags, int kind, int theirSlotForMe, int mySlotForThem);
int mvdPeerSignal(mvdContext* ctx, unsigned int value);
int mvdPeerWait(mvdContext* ctx, unsigned int value);
int mvdPeerBarrier(mvdContext* ctx, unsigned int value);

/* Stand-alone resident convolution with the engine used by the loop: out = conv(ext(src), kernel)
 * restricted to the volume, src/out dim (x,y,z) host arrays.  extMode 0 zero, 1 constant
 * extValue, 2 mirror-single.  Used by the tests and by the kernel preparation. */
int mvdConvolve(int device, const float* src, const int dim[3], const float* kernel, const int kernelDim[3], int extMode,
                float extValue, float* out);

/* Same with device-resident src/out (dSrc != dOut) on the given stream (cudaStream_t as void*,
 * NULL = legacy default stream); the kernel stays a host array.  Synchronises before returning. */
int mvdConvolveDevice(int device, const float* dSrc, const int dim[3], const float* kernel, const int kernelDim[3], int extMode,
                      float extValue, float* dOut, void* cudaStream);

/* Tile plan chosen for (dim, kernelDim): writes up to maxTiles records of 12 ints
 * (window size b[3], window origin o[3], first output voxel e[3], output count n[3]) and returns
 * the tile count, or -1. */
int mvdPlanTiles(const int dim[3], const int kernelDim[3], int* out, int maxTiles);

/* Block tiling of the reference, bit-exact (BlockGeneratorFixedSizePrecise.java:25-101).
 * Writes per block 5*numDimensions values: blockSize, offset, effectiveSize, effectiveOffset,
 * effectiveLocalOffset.  Returns the number of blocks, 0 if blockSize is smaller than the kernel
 * (the reference returns null, :38-42), or -1 if maxBlocks is too small. */
long long mvdDivideIntoBlocks(const long long* imgSize, const long long* kernelSize, const long long* blockSize,
                              int numDimensions, long long* out, long long maxBlocks);

/* Counters for bench.py: kernels launched by this library in the calling process. */
long long mvdKernelLaunchCount(void);

/* Allocation-free separable convolution.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  out = the three in-place 1-D passes of convolve_7 / 15 / 31 (same kernels, same arithmetic),
 * with DEVICE arrays dSrc -> dOut (dSrc != dOut, x fastest, dim = {x, y, z}); a NULL kernel leaves that axis alone;
 * kernelX/Y/Z are HOST arrays of kernelLen <= 31 taps; outofbounds 0 zero, 1 outofboundsvalue, 2 extend last pixel.
 * The call launches two kernels on cudaStream (a cudaStream_t, NULL = legacy default stream) and returns without
 * synchronising; it calls no cudaMalloc / cudaFree and no allocating library routine.  The intermediate lives in a
 * static device array of MVD_NOALLOC_MAX_VOXELS floats (32 MiB, held by every GPU that loads the library until the
 * process exits): volumes with more voxels are rejected with an error (mvdGetLastError).  Every call shares that
 * array -- calls that overlap on the GPU (other streams, other host threads on the same device) race; serialise them.
 * The first call on a device resolves the array's device address (cudaGetSymbolAddress): make it before a capture. */
#define MVD_NOALLOC_MAX_VOXELS 8388608 /* 8 Mi voxels, e.g. 256 x 256 x 128 */
int mvdSeparableConvolveNoAlloc(int device, const float* dSrc, float* dOut, const int dim[3], const float* kernelX,
                                const float* kernelY, const float* kernelZ, int kernelLen, int outofbounds,
                                float outofboundsvalue, void* cudaStream);

/* Process-wide switches between kernel generations, for A/B measurements and tests (defaults: the faster one).
 *   "z_tma"   1 = z pass with TMA bulk-tensor loads / stores (z_tma_kernels.cu), 0 = register-loading z pass (default)
 *   "inv_ring" inverse xy pass fed by bulk-tensor loads into a shared-memory ring: 2 = for the quotient / store epilogues
 *             (default; the update epilogue keeps the register-loading kernel, which is faster there), 1 = always, 0 = never
 *   "z_warp"  z pass with warp-local lines behind swizzled bulk-tensor loads / stores (z_warp_kernels.cu): 4 = lines on part
 *             of a warp, two FFT passes, three-buffer ring (z_part_kernel) for 512- and 256-point windows (default),
 *             3 = the same for 512-point windows only, 1 = the first generation (one line per warp, three passes) for
 *             512-point windows, 2 = also for 256-point ones, 0 = register-loading z pass; set before views are added
 * Returns the previous value, or -1 for an unknown name. */
int mvdSetTuning(const char* name, int value);

#ifdef __cplusplus
}
#endif
#endif /* MVDECON_H */
