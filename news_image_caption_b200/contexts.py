"""Context assembly between the encoders and the decoder call, on the B200 kernels.

Mirrors the tail of `TransformerFacesObjectModel._forward` (ttl/tell/models/transformer_faces_objects.py:
336-395; 2-source variant: transformer_flattened.py `_forward`): image grid flattening, RoBERTa hidden-layer
mix (`weigh_bert`), article padding mask, NaN-padded face / object embeddings -> masks, everything transposed
to the decoder's [S, B, dim] layout.  The encoders themselves (ResNet-152, RoBERTa) are out of scope.
"""
import torch

from . import ops


def assemble_contexts(X_image, article_ids, article_hiddens, bert_weight=None, face_embeds=None, obj_embeds=None,
                      padding_idx=1, dtype=torch.bfloat16):
    """X_image [B,C,H,W] (ResNet feature map, as the reference permutes it at :337-341) or [B,P,C];
    article_ids [B,S]; article_hiddens: list of [B,S,E] (all RoBERTa hidden layers) -- mixed with
    softmax(bert_weight) when bert_weight is given, else the last one is used (:354-369);
    face_embeds [B,n,512] / obj_embeds [B,n,2048] with NaN rows = padding, or None (2-source model)."""
    if X_image.dim() == 4:
        B, Cc, H, W = X_image.shape
        X_image = X_image.permute(0, 2, 3, 1).reshape(B, H * W, Cc)
    B, P, _ = X_image.shape
    ctx = {'image': X_image.transpose(0, 1).to('cuda', dtype).contiguous(),
           'image_mask': torch.zeros(B, P, dtype=torch.bool, device='cuda'),
           'article_mask': (article_ids == padding_idx).to('cuda')}
    if bert_weight is not None:
        ctx['article'] = ops.context_mix(list(article_hiddens), bert_weight, out_dtype=dtype)
    else:
        ctx['article'] = article_hiddens[-1].transpose(0, 1).to('cuda', dtype).contiguous()
    if face_embeds is not None:
        ctx['faces'], ctx['faces_mask'] = ops.nan_mask(face_embeds, out_dtype=dtype)
    if obj_embeds is not None:
        ctx['obj'], ctx['obj_mask'] = ops.nan_mask(obj_embeds, out_dtype=dtype)
    return ctx
