#!/usr/bin/env python3
"""TEST INFRASTRUCTURE ONLY.  Prints one reference source file with the dump / GPU hooks inserted, for the compiler to read.

    make_overlay.py <reference root> <file name>  > <temporary translation unit>

Reads <ref>/Source/Lib/Codec/<file name> (me_process.c, global_me.c, src_ops_process.c, pd_process.c, temporal_filtering.c)
and inserts the calls into oracle/enc_overlay/me_dump.c / me_gpu.c listed in HOOKS below.  oracle/Makefile compiles the output
and deletes it in the same recipe: no copy of a reference source stays on disk, let alone in the repository -- only the object
file under oracle/_ref/enc/.  Every anchor must match exactly once, else the script fails: a reference update that moves the
code is noticed instead of silently hooking nothing.
"""
import pathlib
import sys

DECLS = (
    "\n/* svtme dump hooks (oracle/enc_overlay/me_dump.c) */\n"
    "struct SequenceControlSet; struct PictureParentControlSet; struct MeContext; struct EbPictureBufferDesc;\n"
    "void svtme_dump_pre_sb(struct PictureParentControlSet *pcs, unsigned b64);\n"
    "void svtme_dump_gm_level(struct PictureParentControlSet *pcs, unsigned char level);\n"
    "void svtme_dump_tpl(struct SequenceControlSet *scs, struct PictureParentControlSet *pcs);\n"
    "void svtme_gpu_tpl_prepare(struct SequenceControlSet *scs, struct PictureParentControlSet *pcs);\n"
    "int svtme_gpu_tpl_block(struct PictureParentControlSet *pcs, unsigned x, unsigned y, unsigned char *best_mode, int *best_rf_idx,\n"
    "                        void *final_best_mv, unsigned long long *best_ref_poc);\n"
    "int svtme_gpu_me_picture(struct SequenceControlSet *scs, struct PictureParentControlSet *pcs, struct MeContext *ctx,\n"
    "                         struct EbPictureBufferDesc *input_pic, unsigned segment_index);\n"
    "int svtme_gpu_dg_picture(struct PictureParentControlSet *ppcs, unsigned seg_idx);\n"
    "void svtme_dump_dg(struct PictureParentControlSet *src_pcs);\n"
    "int svtme_gpu_tf_b64(struct PictureParentControlSet *centre_pcs, struct PictureParentControlSet *ref_pcs, struct MeContext *ctx,\n"
    "                     unsigned b64_index);\n"
    "void svtme_dump_tf_b64(struct PictureParentControlSet *centre_pcs, struct PictureParentControlSet *ref_pcs, struct MeContext *ctx,\n"
    "                       unsigned b64_index);\n"
    "void svtme_dump_picture(struct SequenceControlSet *scs, struct PictureParentControlSet *pcs, struct MeContext *ctx,\n"
    "                        struct EbPictureBufferDesc *input_pic);\n"
)

# (file, anchor, text, where): `text` goes before/after the single line that contains `anchor`
HOOKS = [
    # me_process.c:173 -- in front of the 64x64 block loop of a TASK_PAME segment: with SVTME_GPU_ME=1 the whole picture is
    # searched by libsvtme_cuda.so through the encoder-side shim (segment 0 does it, the other segments have nothing left to
    # do), and the loop below runs over an empty range
    ("me_process.c", "// 64x64 Block Loop",
     "if (svtme_gpu_me_picture(scs, pcs, me_context_ptr->me_ctx, input_pic, segment_index)) y_b64_end_index = y_b64_start_index;\n",
     "before"),
    # me_process.c:263 -- before the per-block search
    ("me_process.c", "svt_aom_motion_estimation_b64(pcs,", "svtme_dump_pre_sb(pcs, b64_index);\n", "before"),
    # me_process.c:282 -- the memset of the `else` branch is the last statement of the last-SB hook; the dump goes after
    # the closing brace of that else, i.e. two lines further down
    ("me_process.c", "memset(pcs->is_global_motion, false, MAX_NUM_OF_REF_PIC_LIST * REF_LIST_MAX_DEPTH);",
     "svtme_dump_picture(scs, pcs, me_context_ptr->me_ctx, input_pic);\n", "after_closing_brace"),
    # global_me.c:206 -- after the bypass test, before the level is used
    ("global_me.c", "if (global_motion_estimation_level) {", "svtme_dump_gm_level(pcs, global_motion_estimation_level);\n",
     "before"),
    # pd_process.c:517 -- in front of the block loop of a TASK_DG_DETECTOR_HME segment (dg_detector_hme_level0): with
    # SVTME_GPU_ME=1 segment 0 has the whole picture searched by svtme_dg_detector_hme_level0 and the loop runs over nothing
    ("pd_process.c", "for (uint32_t y_b64_idx = y_b64_start_idx; y_b64_idx < y_b64_end_idx; ++y_b64_idx) {",
     "if (svtme_gpu_dg_picture(ppcs, seg_idx)) y_b64_end_idx = y_b64_start_idx;\n", "before"),
    # pd_process.c:624 -- early_hme has every segment's metrics of one (source, reference) pair
    ("pd_process.c", "svt_block_on_semaphore(src_pcs->dg_detector->frame_done_sem);", "svtme_dump_dg(src_pcs);\n", "after"),
    # temporal_filtering.c:3169 -- the ME call of produce_temporally_filtered_pic becomes conditional: with SVTME_GPU_ME=1 the
    # block's outputs come from the whole-picture ME_MCTF search of the (centre, neighbour) pair (svtme_shim_tf_b64) ...
    ("temporal_filtering.c", "svt_aom_motion_estimation_b64(centre_pcs,",
     "if (!svtme_gpu_tf_b64(centre_pcs, pcs_list[frame_index], ctx, (uint32_t)blk_row * blk_cols + blk_col))\n", "before"),
    # ... and :3174, after the call: what the filter is about to read from the MeContext
    ("temporal_filtering.c", "input_picture_ptr_central); // source picture",
     "svtme_dump_tf_b64(centre_pcs, pcs_list[frame_index], ctx, (uint32_t)blk_row * blk_cols + blk_col);\n", "after"),
    # src_ops_process.c:1378 -- before the picture's dispenser tasks are posted: with SVTME_GPU_ME=1 the source-based decisions of
    # all its TPL blocks come from the GPU (svtme_shim_tpl_picture) ...
    ("src_ops_process.c", "pcs->tpl_disp_coded_sb_count = 0;", "svtme_gpu_tpl_prepare(scs, pcs);\n", "after"),
    # ... and src_ops_process.c:612-887, the intra + inter source search of one block, is skipped when its decision is there
    ("src_ops_process.c", "            if (!disable_intra_pred) {",
     "if (!svtme_gpu_tpl_block(pcs, mb_origin_x, mb_origin_y, &best_mode, &best_rf_idx, &final_best_mv, (unsigned long long *)&best_ref_poc)) {\n",
     "before"),
    ("src_ops_process.c", "                best_mode = NEWMV;", "}\n", "after"),
    # src_ops_process.c:1396 -- every SB of the picture went through tpl_mc_flow_dispenser_sb_generic
    ("src_ops_process.c", "svt_block_on_semaphore(pcs->tpl_disp_done_semaphore);", "svtme_dump_tpl(scs, pcs);\n", "after"),
]


def patch(name, text):
    lines = text.splitlines(keepends=True)
    for fname, anchor, ins, where in HOOKS:
        if fname != name:
            continue
        hits = [i for i, l in enumerate(lines) if anchor in l]
        if len(hits) != 1:
            sys.exit(f"{name}: anchor {anchor!r} matched {len(hits)} times, expected 1")
        i = hits[0]
        if where == "before":
            lines.insert(i, ins)
        elif where == "after":
            lines.insert(i + 1, ins)
        elif where == "after_closing_brace":
            j = i + 1
            while lines[j].strip() != "}":
                j += 1
            lines.insert(j + 1, ins)
        else:
            sys.exit(f"bad hook placement {where}")
    # declarations after the last #include
    last_inc = max(i for i, l in enumerate(lines) if l.startswith("#include"))
    lines.insert(last_inc + 1, DECLS)
    return "".join(lines)


def main():
    ref, name = pathlib.Path(sys.argv[1]), sys.argv[2]
    sys.stdout.write(patch(name, (ref / "Source/Lib/Codec" / name).read_text()))


if __name__ == "__main__":
    main()
