"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): restatement of the reference's greedy search.

Follows rnnt/model.py:74-111 (``Transducer.recognize``; rnnt/search.py:5-40 holds the same loop):
per utterance, ``dec_state, hidden = decoder(zero_token)`` (model.py:86); for every frame
``t < length`` (model.py:88) the joint is evaluated on the two 1-D vectors (model.py:89, the
equal-rank branch model.py:28-30), softmax + argmax (model.py:90-92); a non-zero prediction is
appended and fed to the decoder with the carried hidden state (model.py:94-101). At most one
symbol per frame. Parity is unpinned by the reference (it holds no decoding test or fixture)."""
import torch
import torch.nn.functional as F


@torch.no_grad()
def greedy_search(enc_states, inputs_length, decoder, joint):
    """enc_states (B,T,E); returns a list of B token lists."""
    results = []
    dev = enc_states.device
    zero_token = torch.zeros(1, 1, dtype=torch.long, device=dev)
    for i in range(enc_states.shape[0]):
        enc_state = enc_states[i]
        token_list = []
        dec_state, hidden = decoder(zero_token)
        for t in range(int(inputs_length[i])):
            x = torch.cat((enc_state[t].view(-1), dec_state.view(-1)), dim=-1)
            logits = joint.project_layer(torch.tanh(joint.forward_layer(x)))
            pred = int(torch.argmax(F.softmax(logits, dim=0), dim=0))
            if pred != 0:
                token_list.append(pred)
                dec_state, hidden = decoder(torch.full((1, 1), pred, dtype=torch.long, device=dev), hidden=hidden)
        results.append(token_list)
    return results
