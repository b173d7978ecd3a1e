"""Oracle (test infrastructure): restatement of the multimodal splice, the part of
``prepare_inputs_labels_for_multimodal`` after the event features exist
(``model/EventChatModel.py:355-465``): drop masked tokens, embed the text chunks,
put the 582 event rows where each ``EVENT_TOKEN_INDEX`` (-200) stood, truncate to 2048,
right-pad, and rebuild labels / attention mask / position ids with the reference's
``None`` conventions.  See ``oracle/__init__.py`` for the usage rule.
"""
import torch

IGNORE_INDEX = -100          # dataset/constants.py:7
EVENT_TOKEN_INDEX = -200     # dataset/constants.py:8
MAX_LEN = 2048               # model/EventChatModel.py:415


def splice(embed_table, input_ids, position_ids, attention_mask, labels, event_features):
    """Returns (position_ids, attention_mask, inputs_embeds, labels) exactly as the reference returns
    elements 1, 2, 4, 5 of its 6-tuple (element 0 and 3 are None / past_key_values)."""
    _labels, _position_ids, _attention_mask = labels, position_ids, attention_mask
    if attention_mask is None:
        attention_mask = torch.ones_like(input_ids, dtype=torch.bool)
    else:
        attention_mask = attention_mask.bool()
    if position_ids is None:
        position_ids = torch.arange(0, input_ids.shape[1], dtype=torch.long)
    if labels is None:
        labels = torch.full_like(input_ids, IGNORE_INDEX)
    ids = [i[m] for i, m in zip(input_ids, attention_mask)]              # :372
    labs = [l[m] for l, m in zip(labels, attention_mask)]                # :373
    new_embeds, new_labels, cur = [], [], 0
    for b, cur_ids in enumerate(ids):
        n_ev = int((cur_ids == EVENT_TOKEN_INDEX).sum())
        if n_ev == 0:                                                    # :380-387
            new_embeds.append(torch.cat([embed_table[cur_ids], event_features[cur][0:0]], dim=0))
            new_labels.append(labs[b])
            cur += 1
            continue
        idx = [-1] + torch.where(cur_ids == EVENT_TOKEN_INDEX)[0].tolist() + [cur_ids.shape[0]]   # :389
        e_parts, l_parts = [], []
        for i in range(len(idx) - 1):
            chunk = cur_ids[idx[i] + 1: idx[i + 1]]
            e_parts.append(embed_table[chunk])
            l_parts.append(labs[b][idx[i] + 1: idx[i + 1]])
            if i < n_ev:                                                 # :405-409
                f = event_features[cur]
                cur += 1
                e_parts.append(f)
                l_parts.append(torch.full((f.shape[0],), IGNORE_INDEX, dtype=labs[b].dtype))
        new_embeds.append(torch.cat(e_parts))
        new_labels.append(torch.cat(l_parts))
    new_embeds = [x[:MAX_LEN] for x in new_embeds]                       # :416-418
    new_labels = [x[:MAX_LEN] for x in new_labels]
    max_len = max(x.shape[0] for x in new_embeds)
    B = len(new_embeds)
    out = torch.zeros((B, max_len, embed_table.shape[1]), dtype=embed_table.dtype)
    lab = torch.full((B, max_len), IGNORE_INDEX, dtype=new_labels[0].dtype)
    am = torch.zeros((B, max_len), dtype=attention_mask.dtype)
    pos = torch.zeros((B, max_len), dtype=position_ids.dtype)
    for i, (e, l) in enumerate(zip(new_embeds, new_labels)):             # right padding, :428-447
        n = e.shape[0]
        out[i, :n] = e
        if n > 0:
            lab[i, :n] = l
            am[i, :n] = True
            pos[i, :n] = torch.arange(0, n, dtype=pos.dtype)
    lab = None if _labels is None else lab
    am = None if _attention_mask is None else am.to(dtype=_attention_mask.dtype)
    pos = None if _position_ids is None else pos
    return pos, am, out, lab
