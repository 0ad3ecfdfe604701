package tensorflow.serving;

import java.lang.foreign.Arena;
import java.lang.foreign.FunctionDescriptor;
import java.lang.foreign.Linker;
import java.lang.foreign.MemorySegment;
import java.lang.foreign.SymbolLookup;
import java.lang.invoke.MethodHandle;
import java.util.concurrent.CompletableFuture;
import java.util.concurrent.ExecutorService;
import java.util.concurrent.Executors;

import static java.lang.foreign.ValueLayout.ADDRESS;
import static java.lang.foreign.ValueLayout.JAVA_FLOAT;
import static java.lang.foreign.ValueLayout.JAVA_INT;
import static java.lang.foreign.ValueLayout.JAVA_LONG;

/**
 * Panama (FFM, JDK 22+) binding of include/ctr_engine.h for the reference client.
 *
 * SOURCE ONLY: the build image has no JDK, so this file is not compiled or run there; the
 * tested boundary is the C ABI it calls (ctypes tests and csrc/dcn_client_host.cpp).
 *
 * It keeps the shape of DCNClient's request path: flat id / weight arrays in, flat float
 * array out, one CompletableFuture per request on a fixed pool (DCNClient.java:42, :146-164).
 * What changes is below the future: no protobuf marshalling and no RPC (DCNClient.java:91-115);
 * the engine splits the candidates over its GPUs by the DCNClient.java:46-55 rule and merges
 * the scores in order.
 */
public final class CtrEngine implements AutoCloseable {
    private static final Linker LINKER = Linker.nativeLinker();
    private final Arena arena = Arena.ofShared();
    private final MethodHandle create, destroy, load, predictAsync, wait_, lastError;
    private final MemorySegment engine;
    private final int fieldNum;
    private final ExecutorService threadPool = Executors.newFixedThreadPool(16);      // DCNClient.java:42

    /** One "channel" per GPU (getChannels, DCNClient.java:118-125). */
    public CtrEngine(String libPath, String weightFile, int[] deviceIds, int maxCandidates, int fieldNum) throws Throwable {
        SymbolLookup lib = SymbolLookup.libraryLookup(libPath, arena);
        create = LINKER.downcallHandle(lib.find("ctr_engine_create").orElseThrow(), FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS));
        destroy = LINKER.downcallHandle(lib.find("ctr_engine_destroy").orElseThrow(), FunctionDescriptor.of(JAVA_INT, ADDRESS));
        load = LINKER.downcallHandle(lib.find("ctr_engine_load_weights").orElseThrow(), FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS));
        predictAsync = LINKER.downcallHandle(lib.find("ctr_predict_async").orElseThrow(),
                FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS, JAVA_LONG, ADDRESS, ADDRESS, ADDRESS, ADDRESS));
        wait_ = LINKER.downcallHandle(lib.find("ctr_wait").orElseThrow(), FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_LONG, JAVA_INT));
        lastError = LINKER.downcallHandle(lib.find("ctr_last_error").orElseThrow(), FunctionDescriptor.of(ADDRESS));
        this.fieldNum = fieldNum;

        // ctr_config: int32 struct_size, n_devices, device_ids[8], max_candidates, workers_per_device, shard_mode, min_shard_candidates, host_pack, reserved[6]
        final int cfgBytes = 4 * (2 + 8 + 3 + 8);
        MemorySegment cfg = arena.allocate(cfgBytes, 4);
        cfg.fill((byte) 0);
        cfg.set(JAVA_INT, 0, cfgBytes);
        cfg.set(JAVA_INT, 4, deviceIds.length);
        for (int i = 0; i < deviceIds.length; i++) cfg.set(JAVA_INT, 8 + 4L * i, deviceIds[i]);
        cfg.set(JAVA_INT, 40, maxCandidates);
        cfg.set(JAVA_INT, 44, 4);
        MemorySegment out = arena.allocate(ADDRESS);
        check((int) create.invokeExact(cfg, out));
        engine = out.get(ADDRESS, 0);
        check((int) load.invokeExact(engine, arena.allocateFrom(weightFile)));
    }

    /** processARequest minus the sort (DCNClient.java:137-164): scores[j] belongs to candidate j. */
    public CompletableFuture<float[]> predict(long[] featureIds, float[] featureWts) {
        return CompletableFuture.supplyAsync(() -> {
            int n = featureIds.length / fieldNum;                                       // DCNClient.java:97
            if (n * fieldNum != featureIds.length || featureWts.length != featureIds.length)
                throw new IllegalArgumentException("declared shape [" + n + "," + fieldNum + "] disagrees with the value count");
            try (Arena a = Arena.ofConfined()) {
                MemorySegment ids = a.allocateFrom(JAVA_LONG, featureIds);
                MemorySegment wts = a.allocateFrom(JAVA_FLOAT, featureWts);
                MemorySegment scores = a.allocate(JAVA_FLOAT, Math.max(n, 1));
                MemorySegment ticket = a.allocate(JAVA_LONG);
                check((int) predictAsync.invokeExact(engine, ids, wts, (long) n, scores, MemorySegment.NULL, MemorySegment.NULL, ticket));
                check((int) wait_.invokeExact(engine, ticket.get(JAVA_LONG, 0), -1));   // join, DCNClient.java:158-159
                return scores.asSlice(0, 4L * n).toArray(JAVA_FLOAT);
            } catch (RuntimeException ex) {
                throw ex;
            } catch (Throwable t) {
                throw new RuntimeException(t);                                          // surfaces as CompletionException at join()
            }
        }, threadPool);
    }

    private void check(int rc) throws Throwable {
        if (rc != 0) {
            MemorySegment msg = ((MemorySegment) lastError.invokeExact()).reinterpret(4096);
            throw new IllegalStateException("ctr engine error " + rc + ": " + msg.getString(0));
        }
    }

    /** shutDownChannels (DCNClient.java:127-135) -- but really releases the devices. */
    @Override
    public void close() {
        threadPool.shutdown();
        try {
            int rc = (int) destroy.invokeExact(engine);
        } catch (Throwable ignored) {
        }
        arena.close();
    }
}
