/*
 * JNI shim between io.trino.operator.gpu.PrestoGpuNative and libprestogpu (include/prestogpu.h).
 * NOT BUILT by __graft_entry__.build(): this image has no JDK (no jni.h). Build where one exists with
 *   gcc -shared -fPIC -I$JAVA_HOME/include -I$JAVA_HOME/include/linux -Iinclude jni/prestogpu_jni.c \
 *       -Lpresto_b200 -lprestogpu -o libprestogpu_jni.so
 * Every native declared in java/io/trino/operator/gpu/PrestoGpuNative.java is defined here; the CPU test-suite compiles this
 * file against a stand-in jni.h (tests/emu/jni_stub) and checks exactly that (tests/test_jni_shim.py).
 */
#include <jni.h>
#include <string.h>

#include "prestogpu.h"

/* io.trino.spi.StandardErrorCode names, indexed by pgpu_status */
static const char* const kErrorCode[] = {"OK", "NOT_SUPPORTED", "GENERIC_INTERNAL_ERROR", "NUMERIC_VALUE_OUT_OF_RANGE", "DIVISION_BY_ZERO",
                                         "GENERIC_INSUFFICIENT_RESOURCES", "COMPILER_ERROR", "GENERIC_INTERNAL_ERROR", "GENERIC_INTERNAL_ERROR",
                                         "INVALID_CAST_ARGUMENT"};

static void throwTrino(JNIEnv* env, pgpu_status st, const char* msg) {
    /* new TrinoException(StandardErrorCode.valueOf(name), msg) */
    if ((int)st < 0 || (int)st >= (int)(sizeof kErrorCode / sizeof kErrorCode[0])) st = PGPU_CUDA_ERROR; /* unknown status: GENERIC_INTERNAL_ERROR */
    jclass codes = (*env)->FindClass(env, "io/trino/spi/StandardErrorCode");
    jmethodID valueOf = (*env)->GetStaticMethodID(env, codes, "valueOf", "(Ljava/lang/String;)Lio/trino/spi/StandardErrorCode;");
    jobject code = (*env)->CallStaticObjectMethod(env, codes, valueOf, (*env)->NewStringUTF(env, kErrorCode[st]));
    jclass ex = (*env)->FindClass(env, "io/trino/spi/TrinoException");
    jmethodID ctor = (*env)->GetMethodID(env, ex, "<init>", "(Lio/trino/spi/ErrorCodeSupplier;Ljava/lang/String;)V");
    (*env)->Throw(env, (jthrowable)(*env)->NewObject(env, ex, ctor, code, (*env)->NewStringUTF(env, msg)));
}

JNIEXPORT jlong JNICALL Java_io_trino_operator_gpu_PrestoGpuNative_create(JNIEnv* env, jclass cls, jbyteArray plan, jint device) {
    jsize n = (*env)->GetArrayLength(env, plan);
    jbyte* text = (*env)->GetByteArrayElements(env, plan, NULL);
    pgpu_pipeline* p = NULL;
    pgpu_status st = pgpu_pipeline_create((const uint8_t*)text, (size_t)n, device, &p);
    (*env)->ReleaseByteArrayElements(env, plan, text, JNI_ABORT);
    if (st != PGPU_OK) {
        throwTrino(env, st, pgpu_last_global_error());
        return 0;
    }
    return (jlong)(intptr_t)p;
}

/*
 * addPage: every Java heap array is pinned with GetPrimitiveArrayCritical only for the duration of pgpu_add_page, which
 * copies it into the pipeline's pinned staging ring (flags = 0) before returning — Trino's blocks are never retained and no
 * JNI call is made while arrays are pinned. Whatever pgpu_add_page could WAIT for (compiling the kernel variant of a new page
 * signature, a staging slot whose previous kernel is still running) is done by pgpu_reserve_input BEFORE the critical region,
 * so the region contains memory copies only and cannot stall the collector behind the device.
 */
#define MAX_CH 64
JNIEXPORT void JNICALL Java_io_trino_operator_gpu_PrestoGpuNative_addPage(JNIEnv* env, jclass cls, jlong handle, jint positionCount, jintArray encoding,
        jintArray typeKind, jintArray precision, jintArray scale, jobjectArray values, jintArray valuesOffset, jobjectArray valueIsNull,
        jobjectArray offsets, jobjectArray ids, jintArray idsOffset, jobjectArray dictValues, jobjectArray dictIsNull, jobjectArray dictOffsets,
        jintArray dictPositionCount) {
    pgpu_pipeline* p = (pgpu_pipeline*)(intptr_t)handle;
    jsize nc = (*env)->GetArrayLength(env, encoding);
    pgpu_column cols[MAX_CH], dicts[MAX_CH];
    jarray pinned[MAX_CH * 8];
    void* pinnedPtr[MAX_CH * 8];
    int npinned = 0;
    jint enc[MAX_CH], kind[MAX_CH], prec[MAX_CH], sc[MAX_CH], voff[MAX_CH], ioff[MAX_CH], dcount[MAX_CH];
    if (nc > MAX_CH) { throwTrino(env, PGPU_NOT_SUPPORTED, "more than 64 channels"); return; }
    (*env)->GetIntArrayRegion(env, encoding, 0, nc, enc);
    (*env)->GetIntArrayRegion(env, typeKind, 0, nc, kind);
    (*env)->GetIntArrayRegion(env, precision, 0, nc, prec);
    (*env)->GetIntArrayRegion(env, scale, 0, nc, sc);
    (*env)->GetIntArrayRegion(env, valuesOffset, 0, nc, voff);
    (*env)->GetIntArrayRegion(env, idsOffset, 0, nc, ioff);
    (*env)->GetIntArrayRegion(env, dictPositionCount, 0, nc, dcount);
    /* collect the array references first (JNI calls are not allowed inside the critical region) */
    jarray aValues[MAX_CH], aNulls[MAX_CH], aOffsets[MAX_CH], aIds[MAX_CH], aDV[MAX_CH], aDN[MAX_CH], aDO[MAX_CH];
    for (jsize c = 0; c < nc; c++) {
        aValues[c] = (jarray)(*env)->GetObjectArrayElement(env, values, c);
        aNulls[c] = (jarray)(*env)->GetObjectArrayElement(env, valueIsNull, c);
        aOffsets[c] = (jarray)(*env)->GetObjectArrayElement(env, offsets, c);
        aIds[c] = (jarray)(*env)->GetObjectArrayElement(env, ids, c);
        aDV[c] = (jarray)(*env)->GetObjectArrayElement(env, dictValues, c);
        aDN[c] = (jarray)(*env)->GetObjectArrayElement(env, dictIsNull, c);
        aDO[c] = (jarray)(*env)->GetObjectArrayElement(env, dictOffsets, c);
    }
    {
        int32_t sigEnc[MAX_CH];
        uint8_t sigNull[MAX_CH];
        for (jsize c = 0; c < nc; c++) {
            sigEnc[c] = enc[c];
            sigNull[c] = (uint8_t)(enc[c] == PGPU_ENC_PLAIN ? aNulls[c] != NULL : aDN[c] != NULL);
        }
        pgpu_status rs = pgpu_reserve_input(p, sigEnc, sigNull, (int32_t)nc, positionCount);
        if (rs != PGPU_OK) { throwTrino(env, rs, pgpu_last_error(p)); return; }
    }
#define PIN(arr) ((arr) ? (pinned[npinned] = (arr), pinnedPtr[npinned] = (*env)->GetPrimitiveArrayCritical(env, (arr), NULL), pinnedPtr[npinned++]) : NULL)
    static const int width[] = {1, 8, 4, 4, 8, 8, 1, 1};
    for (jsize c = 0; c < nc; c++) {
        pgpu_column* col = &cols[c];
        memset(col, 0, sizeof *col);
        col->encoding = enc[c]; col->type_kind = kind[c]; col->precision = prec[c]; col->scale = sc[c];
        col->position_count = positionCount;
        int w = (kind[c] == PGPU_DECIMAL && prec[c] > 18) ? 16 : width[kind[c]];
        if (enc[c] == PGPU_ENC_PLAIN) {
            char* v = (char*)PIN(aValues[c]);
            int varWidth = kind[c] == PGPU_VARCHAR || kind[c] == PGPU_VARBINARY;
            col->values = v ? v + (varWidth ? 0 : (size_t)voff[c] * w) : NULL;      /* boolean[] is 1 byte per element */
            col->nulls = (const uint8_t*)PIN(aNulls[c]);
            if (col->nulls) col->nulls += voff[c];
            col->offsets = (const int32_t*)PIN(aOffsets[c]);
            if (col->offsets) col->offsets += voff[c];
        } else {
            pgpu_column* d = &dicts[c];
            memset(d, 0, sizeof *d);
            d->encoding = PGPU_ENC_PLAIN; d->type_kind = kind[c]; d->precision = prec[c]; d->scale = sc[c];
            d->position_count = dcount[c];
            d->values = PIN(aDV[c]);
            d->nulls = (const uint8_t*)PIN(aDN[c]);
            d->offsets = (const int32_t*)PIN(aDO[c]);
            col->dictionary = d;
            if (enc[c] == PGPU_ENC_DICTIONARY) col->ids = (const int32_t*)PIN(aIds[c]) + ioff[c];
        }
    }
    pgpu_status st = pgpu_add_page(p, cols, (int32_t)nc, positionCount, 0);
    for (int i = npinned - 1; i >= 0; i--) (*env)->ReleasePrimitiveArrayCritical(env, pinned[i], pinnedPtr[i], JNI_ABORT);
    if (st != PGPU_OK) throwTrino(env, st, pgpu_last_error(p));
}

JNIEXPORT void JNICALL Java_io_trino_operator_gpu_PrestoGpuNative_addSerializedPage(JNIEnv* env, jclass cls, jlong h, jlong address, jint length, jint positionCount) {
    pgpu_pipeline* p = (pgpu_pipeline*)(intptr_t)h;
    pgpu_status st = pgpu_add_serialized_page(p, (const uint8_t*)(intptr_t)address, (size_t)length, positionCount);
    if (st != PGPU_OK) throwTrino(env, st, pgpu_last_error(p));
}

/* a SerializedPage with its PageCodecMarker set and uncompressed size (PagesSerde.deserialize :131-170) */
JNIEXPORT void JNICALL Java_io_trino_operator_gpu_PrestoGpuNative_addSerializedPageMarked(JNIEnv* env, jclass cls, jlong h, jlong address, jint length, jint positionCount,
                                                                                        jint markers, jint uncompressedSize) {
    pgpu_pipeline* p = (pgpu_pipeline*)(intptr_t)h;
    pgpu_status st = pgpu_add_serialized_page_marked(p, (const uint8_t*)(intptr_t)address, (size_t)length, positionCount, markers, uncompressedSize);
    if (st != PGPU_OK) throwTrino(env, st, pgpu_last_error(p));
}

/* addParquetRowGroup: the chunks are direct (off-heap) buffers, so nothing is pinned; the small per-channel arrays are copied. */
JNIEXPORT void JNICALL Java_io_trino_operator_gpu_PrestoGpuNative_addParquetRowGroup(JNIEnv* env, jclass cls, jlong h, jlong rows, jlongArray address,
                                                                                   jlongArray length, jintArray physicalType, jintArray maxDefinitionLevel,
                                                                                   jlongArray valueCount, jintArray typeLength, jintArray codec) {
    pgpu_pipeline* p = (pgpu_pipeline*)(intptr_t)h;
    jsize n = (*env)->GetArrayLength(env, address);
    if (n > MAX_CH) {
        throwTrino(env, PGPU_NOT_SUPPORTED, "more than 64 channels");
        return;
    }
    jlong addr[MAX_CH], len[MAX_CH], count[MAX_CH];
    jint type[MAX_CH], def[MAX_CH], tl[MAX_CH], cc[MAX_CH];
    (*env)->GetLongArrayRegion(env, address, 0, n, addr);
    (*env)->GetLongArrayRegion(env, length, 0, n, len);
    (*env)->GetLongArrayRegion(env, valueCount, 0, n, count);
    (*env)->GetIntArrayRegion(env, physicalType, 0, n, type);
    (*env)->GetIntArrayRegion(env, maxDefinitionLevel, 0, n, def);
    (*env)->GetIntArrayRegion(env, typeLength, 0, n, tl);
    (*env)->GetIntArrayRegion(env, codec, 0, n, cc);
    pgpu_parquet_chunk chunks[MAX_CH];
    for (jsize c = 0; c < n; c++) {
        chunks[c].data = (const uint8_t*)(intptr_t)addr[c];
        chunks[c].len = (size_t)len[c];
        chunks[c].physical_type = type[c];
        chunks[c].max_definition_level = def[c];
        chunks[c].num_values = count[c];
        chunks[c].type_length = tl[c];
        chunks[c].codec = cc[c];  /* CompressionCodecName ordinal as parquet.thrift numbers it: SNAPPY / GZIP / LZ4_RAW chunks are inflated natively */
    }
    pgpu_status st = pgpu_add_parquet_row_group(p, chunks, (int32_t)n, rows);
    if (st != PGPU_OK) throwTrino(env, st, pgpu_last_error(p));
}

JNIEXPORT jboolean JNICALL Java_io_trino_operator_gpu_PrestoGpuNative_needsInput(JNIEnv* env, jclass cls, jlong h) { return pgpu_needs_input((pgpu_pipeline*)(intptr_t)h) != 0; }
JNIEXPORT jboolean JNICALL Java_io_trino_operator_gpu_PrestoGpuNative_isFinished(JNIEnv* env, jclass cls, jlong h) { return pgpu_is_finished((pgpu_pipeline*)(intptr_t)h) != 0; }
JNIEXPORT void JNICALL Java_io_trino_operator_gpu_PrestoGpuNative_finish(JNIEnv* env, jclass cls, jlong h) {
    pgpu_pipeline* p = (pgpu_pipeline*)(intptr_t)h;
    pgpu_status st = pgpu_finish(p);
    if (st != PGPU_OK) throwTrino(env, st, pgpu_last_error(p));
}
JNIEXPORT jboolean JNICALL Java_io_trino_operator_gpu_PrestoGpuNative_isBlocked(JNIEnv* env, jclass cls, jlong h) { return pgpu_is_blocked((pgpu_pipeline*)(intptr_t)h) != 0; }
JNIEXPORT jstring JNICALL Java_io_trino_operator_gpu_PrestoGpuNative_lastError(JNIEnv* env, jclass cls, jlong h) {
    return (*env)->NewStringUTF(env, pgpu_last_error((pgpu_pipeline*)(intptr_t)h));
}

/* outputTypes: (kind, precision, scale) per output channel */
JNIEXPORT jintArray JNICALL Java_io_trino_operator_gpu_PrestoGpuNative_outputTypes(JNIEnv* env, jclass cls, jlong h) {
    pgpu_pipeline* p = (pgpu_pipeline*)(intptr_t)h;
    int32_t n = pgpu_output_channel_count(p);
    jintArray out = (*env)->NewIntArray(env, 3 * n);
    if (!out) return NULL;
    for (int32_t c = 0; c < n; c++) {
        jint t[3];
        int32_t kind = 0, prec = 0, scale = 0;
        pgpu_output_channel_type(p, c, &kind, &prec, &scale);
        t[0] = kind; t[1] = prec; t[2] = scale;
        (*env)->SetIntArrayRegion(env, out, 3 * c, 3, t);
    }
    return out;
}

/*
 * onUnblocked: completes a Java future from the CUDA host-callback thread. The callback thread is not a JVM thread, so it
 * attaches, calls Runnable.run() on the global reference it was given, releases it and detaches.
 */
static JavaVM* g_vm;
static void runRunnable(void* user) {
    JNIEnv* env = NULL;
    if ((*g_vm)->AttachCurrentThreadAsDaemon(g_vm, (void**)&env, NULL) != JNI_OK) return;
    jobject r = (jobject)user;
    jclass rc = (*env)->GetObjectClass(env, r);
    jmethodID run = (*env)->GetMethodID(env, rc, "run", "()V");
    if (run) (*env)->CallVoidMethod(env, r, run);
    (*env)->DeleteGlobalRef(env, r);
    (*g_vm)->DetachCurrentThread(g_vm);
}
JNIEXPORT void JNICALL Java_io_trino_operator_gpu_PrestoGpuNative_onUnblocked(JNIEnv* env, jclass cls, jlong h, jobject runnable) {
    pgpu_pipeline* p = (pgpu_pipeline*)(intptr_t)h;
    if (!g_vm) (*env)->GetJavaVM(env, &g_vm);
    jobject ref = (*env)->NewGlobalRef(env, runnable);
    pgpu_status st = pgpu_on_unblocked(p, runRunnable, ref);
    if (st != PGPU_OK) {
        (*env)->DeleteGlobalRef(env, ref);
        throwTrino(env, st, pgpu_last_error(p));
    }
}
JNIEXPORT void JNICALL Java_io_trino_operator_gpu_PrestoGpuNative_destroy(JNIEnv* env, jclass cls, jlong h) { pgpu_pipeline_destroy((pgpu_pipeline*)(intptr_t)h); }

/*
 * getOutput: null when there is no page now; else Object[1 + 3 * channels] =
 *   [0] Integer positionCount, then per channel: values (long[] / int[] / byte[]; long[2n] for long decimals; byte[] payload for
 *   var-width), boolean[] valueIsNull or null, int[] offsets (var-width) or null.
 * The library-owned pinned buffers are copied into fresh Java arrays (Set<Type>ArrayRegion): output blocks belong to the consumer.
 */
JNIEXPORT jobjectArray JNICALL Java_io_trino_operator_gpu_PrestoGpuNative_getOutput(JNIEnv* env, jclass cls, jlong h) {
    pgpu_pipeline* p = (pgpu_pipeline*)(intptr_t)h;
    pgpu_column cols[MAX_CH];
    int32_t nc = 0;
    int64_t rows = -1;
    pgpu_status st = pgpu_get_output(p, cols, MAX_CH, &nc, &rows);
    if (st != PGPU_OK) {
        throwTrino(env, st, pgpu_last_error(p));
        return NULL;
    }
    if (rows < 0) return NULL;
    jclass objectClass = (*env)->FindClass(env, "java/lang/Object");
    jobjectArray out = (*env)->NewObjectArray(env, 1 + 3 * nc, objectClass, NULL);
    if (!out) return NULL;
    jclass integerClass = (*env)->FindClass(env, "java/lang/Integer");
    jmethodID valueOf = (*env)->GetStaticMethodID(env, integerClass, "valueOf", "(I)Ljava/lang/Integer;");
    (*env)->SetObjectArrayElement(env, out, 0, (*env)->CallStaticObjectMethod(env, integerClass, valueOf, (jint)rows));
    const jsize n = (jsize)rows;
    for (int32_t c = 0; c < nc; c++) {
        const pgpu_column* col = &cols[c];
        jobject values = NULL, offsets = NULL, nulls = NULL;
        const int kind = col->type_kind;
        if (kind == PGPU_VARCHAR || kind == PGPU_VARBINARY) {
            const jsize bytes = n ? col->offsets[n] : 0;
            jbyteArray v = (*env)->NewByteArray(env, bytes);
            jintArray o = (*env)->NewIntArray(env, n + 1);
            if (!v || !o) return NULL;
            if (bytes) (*env)->SetByteArrayRegion(env, v, 0, bytes, (const jbyte*)col->values);
            if (n) (*env)->SetIntArrayRegion(env, o, 0, n + 1, (const jint*)col->offsets);
            values = v;
            offsets = o;
        } else if (kind == PGPU_BOOLEAN) {
            jbyteArray v = (*env)->NewByteArray(env, n);
            if (!v) return NULL;
            if (n) (*env)->SetByteArrayRegion(env, v, 0, n, (const jbyte*)col->values);
            values = v;
        } else if (kind == PGPU_INTEGER || kind == PGPU_DATE) {
            jintArray v = (*env)->NewIntArray(env, n);
            if (!v) return NULL;
            if (n) (*env)->SetIntArrayRegion(env, v, 0, n, (const jint*)col->values);
            values = v;
        } else {
            const jsize words = (kind == PGPU_DECIMAL && col->precision > 18) ? 2 * n : n;  /* Int128ArrayBlock: 2 longs per position */
            jlongArray v = (*env)->NewLongArray(env, words);
            if (!v) return NULL;
            if (words) (*env)->SetLongArrayRegion(env, v, 0, words, (const jlong*)col->values);
            values = v;
        }
        if (col->nulls && n) {
            int any = 0;
            for (jsize i = 0; i < n && !any; i++) any = col->nulls[i] != 0;
            if (any) {
                jbooleanArray b = (*env)->NewBooleanArray(env, n);
                if (!b) return NULL;
                (*env)->SetBooleanArrayRegion(env, b, 0, n, (const jboolean*)col->nulls);
                nulls = b;
            }
        }
        (*env)->SetObjectArrayElement(env, out, 1 + 3 * c, values);
        (*env)->SetObjectArrayElement(env, out, 2 + 3 * c, nulls);
        (*env)->SetObjectArrayElement(env, out, 3 + 3 * c, offsets);
    }
    return out;
}
