/* planes.h -- TEST INFRASTRUCTURE ONLY: picture descriptors shared by the compiled-reference harness and the oracle. */
#ifndef SVTME_ORACLE_PLANES_H
#define SVTME_ORACLE_PLANES_H
#include <stdint.h>
/* A padded 8-bit luma plane laid out like EbPictureBufferDesc (pic_buffer_desc.h:34-75):
 * pixel (0,0) is at buf[org_y*stride + org_x]. */
typedef struct SvtRefPlane {
    const uint8_t *buf;
    uint32_t       stride, width, height, org_x, org_y;
} SvtRefPlane;
typedef struct SvtRefPic {
    SvtRefPlane lvl[3]; /* 0 full, 1 quarter, 2 sixteenth */
    uint64_t    picture_number;
} SvtRefPic;
#endif
