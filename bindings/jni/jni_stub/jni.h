/* Stand-in for the JDK's <jni.h> in an image without a JDK.  The function table has the LAYOUT of the JNI
 * specification (Java SE 11, "JNI Functions", interface function table: 4 reserved slots, GetVersion at 4, FindClass at
 * 6, ThrowNew at 14, GetArrayLength at 171, Get<Prim>ArrayElements at 183.., GetDirectBufferAddress at 230, 234 slots
 * up to GetModule), with the functions bindings/jni/qegpu_jni.c calls declared and every other slot an untyped
 * placeholder -- so the shim compiled against this header calls through the slots a real JVM fills.  Used to
 * syntax-check the shim (tests/test_capi_cpu.py) and to EXECUTE it against a mock JNIEnv (tests/jni_mock.c,
 * tests/test_gpu_jni.py).  Never shipped; a real build uses $JAVA_HOME/include/jni.h. */
#ifndef QE_JNI_STUB_H
#define QE_JNI_STUB_H
#include <stdint.h>
typedef int32_t jint; typedef int64_t jlong; typedef uint8_t jboolean; typedef double jdouble; typedef jint jsize;
typedef void *jobject; typedef jobject jclass; typedef jobject jarray; typedef jobject jthrowable;
typedef jarray jintArray; typedef jarray jlongArray; typedef jarray jbooleanArray;
#define JNIEXPORT __attribute__((visibility("default")))
#define JNICALL
#define JNI_ABORT 2
#define JNI_COMMIT 1
#define JNI_TRUE 1
#define JNI_FALSE 0
struct JNINativeInterface_;
typedef const struct JNINativeInterface_ *JNIEnv;
struct JNINativeInterface_ {
    void *slot0;
    void *slot1;
    void *slot2;
    void *slot3;
    void *slot4;
    void *slot5;
    jclass (*FindClass)(JNIEnv *, const char *);   /* 6 */
    void *slot7;
    void *slot8;
    void *slot9;
    void *slot10;
    void *slot11;
    void *slot12;
    void *slot13;
    jint (*ThrowNew)(JNIEnv *, jclass, const char *);   /* 14 */
    jobject (*ExceptionOccurred)(JNIEnv *);   /* 15 */
    void *slot16;
    void (*ExceptionClear)(JNIEnv *);   /* 17 */
    void *slot18;
    void *slot19;
    void *slot20;
    void *slot21;
    void *slot22;
    void *slot23;
    void *slot24;
    void *slot25;
    void *slot26;
    void *slot27;
    void *slot28;
    void *slot29;
    void *slot30;
    void *slot31;
    void *slot32;
    void *slot33;
    void *slot34;
    void *slot35;
    void *slot36;
    void *slot37;
    void *slot38;
    void *slot39;
    void *slot40;
    void *slot41;
    void *slot42;
    void *slot43;
    void *slot44;
    void *slot45;
    void *slot46;
    void *slot47;
    void *slot48;
    void *slot49;
    void *slot50;
    void *slot51;
    void *slot52;
    void *slot53;
    void *slot54;
    void *slot55;
    void *slot56;
    void *slot57;
    void *slot58;
    void *slot59;
    void *slot60;
    void *slot61;
    void *slot62;
    void *slot63;
    void *slot64;
    void *slot65;
    void *slot66;
    void *slot67;
    void *slot68;
    void *slot69;
    void *slot70;
    void *slot71;
    void *slot72;
    void *slot73;
    void *slot74;
    void *slot75;
    void *slot76;
    void *slot77;
    void *slot78;
    void *slot79;
    void *slot80;
    void *slot81;
    void *slot82;
    void *slot83;
    void *slot84;
    void *slot85;
    void *slot86;
    void *slot87;
    void *slot88;
    void *slot89;
    void *slot90;
    void *slot91;
    void *slot92;
    void *slot93;
    void *slot94;
    void *slot95;
    void *slot96;
    void *slot97;
    void *slot98;
    void *slot99;
    void *slot100;
    void *slot101;
    void *slot102;
    void *slot103;
    void *slot104;
    void *slot105;
    void *slot106;
    void *slot107;
    void *slot108;
    void *slot109;
    void *slot110;
    void *slot111;
    void *slot112;
    void *slot113;
    void *slot114;
    void *slot115;
    void *slot116;
    void *slot117;
    void *slot118;
    void *slot119;
    void *slot120;
    void *slot121;
    void *slot122;
    void *slot123;
    void *slot124;
    void *slot125;
    void *slot126;
    void *slot127;
    void *slot128;
    void *slot129;
    void *slot130;
    void *slot131;
    void *slot132;
    void *slot133;
    void *slot134;
    void *slot135;
    void *slot136;
    void *slot137;
    void *slot138;
    void *slot139;
    void *slot140;
    void *slot141;
    void *slot142;
    void *slot143;
    void *slot144;
    void *slot145;
    void *slot146;
    void *slot147;
    void *slot148;
    void *slot149;
    void *slot150;
    void *slot151;
    void *slot152;
    void *slot153;
    void *slot154;
    void *slot155;
    void *slot156;
    void *slot157;
    void *slot158;
    void *slot159;
    void *slot160;
    void *slot161;
    void *slot162;
    void *slot163;
    void *slot164;
    void *slot165;
    void *slot166;
    void *slot167;
    void *slot168;
    void *slot169;
    void *slot170;
    jsize (*GetArrayLength)(JNIEnv *, jarray);   /* 171 */
    void *slot172;
    void *slot173;
    void *slot174;
    void *slot175;
    void *slot176;
    void *slot177;
    void *slot178;
    void *slot179;
    void *slot180;
    void *slot181;
    void *slot182;
    jboolean *(*GetBooleanArrayElements)(JNIEnv *, jbooleanArray, jboolean *);   /* 183 */
    void *slot184;
    void *slot185;
    void *slot186;
    jint *(*GetIntArrayElements)(JNIEnv *, jintArray, jboolean *);   /* 187 */
    jlong *(*GetLongArrayElements)(JNIEnv *, jlongArray, jboolean *);   /* 188 */
    void *slot189;
    void *slot190;
    void (*ReleaseBooleanArrayElements)(JNIEnv *, jbooleanArray, jboolean *, jint);   /* 191 */
    void *slot192;
    void *slot193;
    void *slot194;
    void (*ReleaseIntArrayElements)(JNIEnv *, jintArray, jint *, jint);   /* 195 */
    void (*ReleaseLongArrayElements)(JNIEnv *, jlongArray, jlong *, jint);   /* 196 */
    void *slot197;
    void *slot198;
    void *slot199;
    void *slot200;
    void *slot201;
    void *slot202;
    void *slot203;
    void *slot204;
    void *slot205;
    void *slot206;
    void (*SetBooleanArrayRegion)(JNIEnv *, jbooleanArray, jsize, jsize, const jboolean *);   /* 207 */
    void *slot208;
    void *slot209;
    void *slot210;
    void *slot211;
    void *slot212;
    void *slot213;
    void *slot214;
    void *slot215;
    void *slot216;
    void *slot217;
    void *slot218;
    void *slot219;
    void *slot220;
    void *slot221;
    void *slot222;
    void *slot223;
    void *slot224;
    void *slot225;
    void *slot226;
    void *slot227;
    jboolean (*ExceptionCheck)(JNIEnv *);   /* 228 */
    void *slot229;
    void *(*GetDirectBufferAddress)(JNIEnv *, jobject);   /* 230 */
    jlong (*GetDirectBufferCapacity)(JNIEnv *, jobject);   /* 231 */
    void *slot232;
    void *slot233;
};
#endif
