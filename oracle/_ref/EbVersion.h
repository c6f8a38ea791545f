#define SVT_AV1_CVS_VERSION "v3.0.2-oracle-build"
